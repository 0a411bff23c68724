"""Deterministic stand-in for the LLM: a seeded population of distinct policy sources.

BASELINE config 2 asks for "1,024 LLM-mutated policy variants"; the GPU box has no network, so the
population is generated here: the parameterised family of the reference's
_create_random_policy (funsearch/funsearch_integration.py:403-431: four uniform floats) plus
constant-jitter / term-dropout mutations of a set of hand-written policy bodies in the style the
reference's LLM prompt asks for (funsearch/safe_execution.py:226-260), each wrapped in the
reference's feasibility prefix and `return max(1, int(score))` suffix (:204-223).
Every source returned translates with policy_compiler.compile_policy.
"""
from __future__ import annotations

import random
import re
from typing import List

from .policy_compiler import PolicyCompileError, compile_policy

PREFIX = '''
def priority_function(pod, node):
    if (pod.cpu_milli > node.cpu_milli_left or
        pod.memory_mib > node.memory_mib_left or
        pod.num_gpu > node.gpu_left):
        return 0

    if pod.num_gpu > 0:
        available_gpus = 0
        for gpu in node.gpus:
            if gpu.gpu_milli_left >= pod.gpu_milli:
                available_gpus += 1
        if available_gpus < pod.num_gpu:
            return 0

    score = 0.0
'''
SUFFIX = '''
    return max(1, int(score))
'''

# Bodies: lists of independent "terms"; a mutation jitters the numeric constants and may drop terms.
BODIES = [
    [
        "    cpu_after = (node.cpu_milli_left - pod.cpu_milli) / node.cpu_milli_total",
        "    mem_after = (node.memory_mib_left - pod.memory_mib) / node.memory_mib_total",
        "    score += (1.0 - cpu_after) * 400.0",
        "    score += (1.0 - mem_after) * 300.0",
        "    score -= abs(cpu_after - mem_after) * 120.0",
        "    if pod.num_gpu > 0:\n        score += (1.0 - (node.gpu_left - pod.num_gpu) / max(1, len(node.gpus))) * 500.0",
    ],
    [
        "    cpu_util = (node.cpu_milli_total - node.cpu_milli_left + pod.cpu_milli) / max(1, node.cpu_milli_total)",
        "    mem_util = (node.memory_mib_total - node.memory_mib_left + pod.memory_mib) / max(1, node.memory_mib_total)",
        "    score += (cpu_util * mem_util) ** 0.5 * 900.0",
        "    score += (1 - abs(cpu_util - mem_util)) ** 2.0 * 250.0",
        "    if pod.num_gpu > 0:\n        fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]\n"
        "        tight = sorted(fits, key=lambda g: g.gpu_milli_left)[:pod.num_gpu]\n"
        "        score += sum(1 - (g.gpu_milli_left - pod.gpu_milli) / g.gpu_milli_total for g in tight) / pod.num_gpu * 700.0",
    ],
    [
        "    free_cpu = node.cpu_milli_left / node.cpu_milli_total",
        "    free_mem = node.memory_mib_left / node.memory_mib_total",
        "    if free_cpu < 0.35:\n        score += (1.0 - free_cpu) * 600.0\n    else:\n        score += (1.0 - free_cpu) * 220.0",
        "    if free_mem < 0.35:\n        score += (1.0 - free_mem) * 450.0\n    else:\n        score += (1.0 - free_mem) * 180.0",
        "    if pod.num_gpu > 0:\n        left = sum(g.gpu_milli_left for g in node.gpus)\n"
        "        score += 800.0 - (left % max(1, pod.gpu_milli)) * 0.4\n"
        "        score -= (max(g.gpu_milli_left for g in node.gpus) - min(g.gpu_milli_left for g in node.gpus)) * 0.06",
        "    if node.cpu_milli_left > pod.cpu_milli * 2 and node.memory_mib_left > pod.memory_mib * 2:\n        score += 35.0",
    ],
    [
        "    slack = min(node.cpu_milli_left - pod.cpu_milli, node.memory_mib_left - pod.memory_mib)",
        "    score += 1500.0 - slack ** 0.6 / max(node.cpu_milli_total, node.memory_mib_total) * 9000.0",
        "    score += (pod.cpu_milli / max(1, node.cpu_milli_left)) * 300.0",
        "    score += (pod.memory_mib / max(1, node.memory_mib_left)) * 200.0",
        "    if pod.num_gpu > 0 and node.gpu_left > 0:\n        score += (pod.num_gpu / node.gpu_left) * 600.0",
    ],
    [
        "    rem_cpu = node.cpu_milli_left - pod.cpu_milli",
        "    rem_mem = node.memory_mib_left - pod.memory_mib",
        "    score = 5000.0 - rem_cpu * 0.01 - rem_mem * 0.002",
        "    if pod.num_gpu > 0:\n        score += 2000.0 - (node.gpu_left - pod.num_gpu) * 240.0",
        "    if pod.num_gpu == 0 and len(node.gpus) > 0:\n        score -= 1500.0",
        "    score += (rem_cpu % max(1, pod.cpu_milli)) / node.cpu_milli_total * -80.0",
    ],
    [
        "    weights = [0.4, 0.35, 0.25]",
        "    a = (node.cpu_milli_left - pod.cpu_milli) / node.cpu_milli_total",
        "    b = (node.memory_mib_left - pod.memory_mib) / node.memory_mib_total",
        "    c = (node.gpu_left - pod.num_gpu) / max(len(node.gpus), 1)",
        "    score = (1 - (a * weights[0] + b * weights[1] + c * weights[2])) * 10000.0",
        "    if 0.3 < a < 0.9:\n        score += 800.0",
        "    score += 300.0 if b < 0.2 else 0.0",
    ],
]


def _jitter_numbers(text: str, rng: random.Random, spread: float) -> str:
    def repl(m: re.Match) -> str:
        tok = m.group(0)
        v = float(tok)
        if v == 0.0:
            return tok
        return repr(round(v * rng.uniform(1.0 - spread, 1.0 + spread), 6))
    # floats with a decimal point only: structural integers (indices, 1, 2) stay as they are
    return re.sub(r"(?<![\w.])\d+\.\d+(?![\w.])", repl, text)


def random_family_policy(rng: random.Random) -> str:
    """Same four-parameter family as the reference's _create_random_policy."""
    cpu_factor = rng.uniform(0.0001, 0.01)
    mem_factor = rng.uniform(0.00001, 0.001)
    gpu_factor = rng.uniform(10.0, 1000.0)
    base_score = rng.uniform(1000.0, 5000.0)
    body = (f"    score = {base_score} + node.cpu_milli_left * {cpu_factor} + node.memory_mib_left * {mem_factor}\n"
            f"    if pod.num_gpu > 0 and node.gpu_left > 0:\n        score += node.gpu_left * {gpu_factor}\n")
    return PREFIX + body + SUFFIX


def mutated_policy(rng: random.Random) -> str:
    terms = list(rng.choice(BODIES))
    defs = [t for t in terms if re.match(r"\s+\w+ = ", t) and "score" not in t.split("=")[0]]
    rest = [t for t in terms if t not in defs]
    if len(rest) > 2 and rng.random() < 0.5:
        del rest[rng.randrange(len(rest))]
    body = "\n".join(defs + rest) + "\n"
    return PREFIX + _jitter_numbers(body, rng, 0.35) + SUFFIX


def policy_population(n: int, seed: int = 0) -> List[str]:
    """n distinct, translatable policy sources (deterministic in `seed`)."""
    rng = random.Random(seed)
    out, seen = [], set()
    guard = 0
    while len(out) < n:
        guard += 1
        if guard > 50 * n + 100:
            raise RuntimeError("could not generate enough distinct policies")
        src = random_family_policy(rng) if rng.random() < 0.25 else mutated_policy(rng)
        if src in seen:
            continue
        try:
            compile_policy(src)
        except PolicyCompileError:
            continue
        seen.add(src)
        out.append(src)
    return out
