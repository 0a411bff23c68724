"""Deterministic stand-in for the LLM: a seeded generation of mutated policy sources.

BASELINE config 2 asks for "1,024 LLM-mutated policy variants"; the GPU box has no network, so the generation is
produced here as SURVEY.md section 8d describes: seeded mutations of the shipped policy sources -- the five of
tests/test_scheduler.py:20-218 incl. the three FunSearch champions and the six baselines of
funsearch/funsearch_integration.py:217-431 (data/base_policies.json) -- plus the parameterised family of the
reference's _create_random_policy (:403-431) and a few template-style bodies written in the style the reference's
LLM prompt asks for (funsearch/safe_execution.py:226-260).

Mutations work on the AST and are of the kinds an LLM produces when it is shown parent policies: constants are
re-tuned (floats and ints), terms are added, dropped or reordered, operators and comparison strictness flip,
a term gets a guard, an expression gets clamped.  Most of them change the *structure* of a policy (its code after
the float literals are lifted out, policy_compiler.py), so a generation holds hundreds of distinct device functions,
not a handful with different constants.  Every source passes the reference's two validators
(`admitted_by_reference_validators` restates safe_execution.py:35-79).
"""
from __future__ import annotations

import ast
import copy
import json
import os
import random
import re
from typing import List, Optional, Sequence

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

# Template-style bodies: lists of independent "terms".
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

# Terms an LLM typically adds: {c} = a coefficient, {e} = an exponent.
TERMS = [
    "(node.cpu_milli_left - pod.cpu_milli) / node.cpu_milli_total * {c}",
    "(node.memory_mib_left - pod.memory_mib) / node.memory_mib_total * {c}",
    "(1.0 - node.cpu_milli_left / node.cpu_milli_total) * {c}",
    "(1.0 - node.memory_mib_left / node.memory_mib_total) * {c}",
    "abs(node.cpu_milli_left / node.cpu_milli_total - node.memory_mib_left / node.memory_mib_total) * {c}",
    "min(node.cpu_milli_left / node.cpu_milli_total, node.memory_mib_left / node.memory_mib_total) * {c}",
    "max(pod.cpu_milli / node.cpu_milli_total, pod.memory_mib / node.memory_mib_total) * {c}",
    "pod.cpu_milli / max(1, node.cpu_milli_left) * {c}",
    "(pod.memory_mib / max(1, node.memory_mib_left)) ** {e} * {c}",
    "((node.cpu_milli_total - node.cpu_milli_left + pod.cpu_milli) / node.cpu_milli_total) ** {e} * {c}",
    "(node.gpu_left - pod.num_gpu) * {c}",
    "len(node.gpus) * {c}",
    "sum(g.gpu_milli_left for g in node.gpus) / max(1, len(node.gpus)) * {c}",
    "len([g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]) * {c}",
    "pod.gpu_milli * pod.num_gpu / max(1, sum(g.gpu_milli_total for g in node.gpus)) * {c}",
    "({c} if node.gpu_left == len(node.gpus) else {c})",
    "(node.cpu_milli_left - pod.cpu_milli) % max(1, pod.cpu_milli) / node.cpu_milli_total * {c}",
    "min(g.gpu_milli_left for g in node.gpus) * {c} if len(node.gpus) > 0 else {c}",
]
GUARDS = [
    "pod.num_gpu > 0", "pod.num_gpu == 0", "node.gpu_left > {i}", "len(node.gpus) > 0",
    "node.cpu_milli_left > pod.cpu_milli * {i}", "node.memory_mib_left > pod.memory_mib * {i}",
    "node.cpu_milli_left / node.cpu_milli_total < {f}", "node.memory_mib_left / node.memory_mib_total > {f}",
    "pod.gpu_milli < {gm}", "pod.cpu_milli > {cm}", "pod.num_gpu > 0 and node.gpu_left >= pod.num_gpu + {i}",
]

FORBIDDEN_PATTERNS = ['import', '__', 'exec', 'eval', 'open', 'file', 'input', 'raw_input', 'compile', 'globals', 'locals',
                      'vars', 'dir', 'hasattr', 'getattr', 'setattr', 'delattr']
ALLOWED_CALL_NAMES = {'abs', 'min', 'max', 'sum', 'len', 'range', 'enumerate', 'int', 'float', 'bool', 'str', 'round', 'sorted',
                      'sqrt', 'log', 'exp', 'pow', 'sin', 'cos', 'tan', 'add', 'sub', 'mul', 'truediv', 'mod'}


def admitted_by_reference_validators(code: str) -> bool:
    """SafeExecutor.validate_code_content + validate_code_structure (funsearch/safe_execution.py:35-79), restated:
    none of the forbidden substrings (case-insensitive), no imports, no dunder attributes, bare-name calls only to the
    whitelisted builtins and math/operator function names."""
    low = code.lower()
    if any(p in low for p in FORBIDDEN_PATTERNS):
        return False
    try:
        tree = ast.parse(code)
    except SyntaxError:
        return False
    for node in ast.walk(tree):
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            return False
        if isinstance(node, ast.Attribute) and node.attr.startswith('__'):
            return False
        if isinstance(node, ast.Call) and isinstance(node.func, ast.Name) and node.func.id not in ALLOWED_CALL_NAMES:
            return False
    return True


def base_policies() -> dict:
    """The shipped policy sources the generation is mutated from (name -> source)."""
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "base_policies.json")) as f:
        return json.load(f)


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


def template_body_policy(rng: random.Random) -> str:
    terms = list(rng.choice(BODIES))
    defs = [t for t in terms if re.match(r"\s+\w+ = ", t) and "score" not in t.split("=")[0]]
    rest = [t for t in terms if t not in defs]
    if len(rest) > 2 and rng.random() < 0.5:
        del rest[rng.randrange(len(rest))]
    body = "\n".join(defs + rest) + "\n"
    return PREFIX + _jitter_numbers(body, rng, 0.35) + SUFFIX


# ---------------------------------------------------------------------------------------- AST mutations
def _coef(rng: random.Random) -> str:
    v = rng.choice([0.05, 0.5, 5.0, 50.0, 500.0]) * rng.uniform(0.4, 2.5) * rng.choice([1, 1, 1, -1])
    return repr(round(v, 3)) if rng.random() < 0.7 else str(int(v * 10) or 3)


def _fill(template: str, rng: random.Random) -> str:
    out = template
    while "{c}" in out:
        out = out.replace("{c}", _coef(rng), 1)
    out = out.replace("{e}", rng.choice(["0.5", "2", "1.5", "2.0", "0.75", "3"]))
    out = out.replace("{i}", str(rng.randint(0, 3)))
    out = out.replace("{f}", repr(round(rng.uniform(0.1, 0.9), 2)))
    out = out.replace("{gm}", str(rng.choice([250, 500, 750, 1000])))
    out = out.replace("{cm}", str(rng.choice([2000, 4000, 8000, 16000])))
    return out


def _prefix_len(fn: ast.FunctionDef) -> int:
    """Leading statements that are the docstring or feasibility guards (`if ...: return 0` blocks): never mutated."""
    k = 0
    for st in fn.body:
        if isinstance(st, ast.Expr) and isinstance(st.value, ast.Constant):
            k += 1
            continue
        if isinstance(st, ast.If) and any(isinstance(n, ast.Return) for n in ast.walk(st)) and not st.orelse:
            k += 1
            continue
        break
    return k


def _score_name(fn: ast.FunctionDef, start: int) -> Optional[str]:
    """The accumulator the policy returns: the name most often updated at function level."""
    cnt: dict = {}
    for st in fn.body[start:]:
        for n in ast.walk(st):
            if isinstance(n, ast.AugAssign) and isinstance(n.target, ast.Name):
                cnt[n.target.id] = cnt.get(n.target.id, 0) + 2
            elif isinstance(n, ast.Assign) and len(n.targets) == 1 and isinstance(n.targets[0], ast.Name):
                cnt[n.targets[0].id] = cnt.get(n.targets[0].id, 0) + 1
    for pref in ("score", "base_score", "final_score"):
        if pref in cnt:
            return pref
    return max(cnt, key=cnt.get) if cnt else None


def _is_update(st: ast.stmt, name: str) -> bool:
    return isinstance(st, ast.AugAssign) and isinstance(st.target, ast.Name) and st.target.id == name \
        and isinstance(st.op, (ast.Add, ast.Sub))


def _only_updates(st: ast.stmt, name: str) -> bool:
    return isinstance(st, ast.If) and not st.orelse and all(_is_update(b, name) for b in st.body)


def _parse_stmt(text: str) -> ast.stmt:
    return ast.parse(text).body[0]


def _parse_expr(text: str) -> ast.expr:
    return ast.parse(text, mode="eval").body


def _template_ending(fn: ast.FunctionDef) -> None:
    """Every policy the reference's generator produces is the template filled in (safe_execution.py:171-224,283-317),
    so it ends with `return max(1, int(score))`: a feasible node never scores 0.  Mutants of the hand-written baselines
    that return a raw number get the same ending (without it a negative term makes feasible nodes score 0 and the pod
    retries at every deletion: 2 M events instead of 20 k)."""
    last = fn.body[-1] if fn.body else None
    if not isinstance(last, ast.Return) or last.value is None:
        return
    v = last.value
    if isinstance(v, ast.Call) and isinstance(v.func, ast.Name) and v.func.id == "max" and len(v.args) == 2 \
            and isinstance(v.args[0], ast.Constant) and v.args[0].value == 1:
        return
    last.value = ast.Call(func=ast.Name(id="max", ctx=ast.Load()),
                          args=[ast.Constant(value=1),
                                ast.Call(func=ast.Name(id="int", ctx=ast.Load()), args=[v], keywords=[])], keywords=[])


def mutate_source(src: str, rng: random.Random, n_mutations: int = 1) -> str:
    """`src` with n_mutations random edits applied (each one of the kinds listed in the module docstring).  The feasibility
    guards at the top of the function are left alone, so a mutant stays a policy that places pods."""
    tree = ast.parse(src)
    fn = [s for s in tree.body if isinstance(s, ast.FunctionDef) and s.name == "priority_function"][-1]
    start = _prefix_len(fn)
    name = _score_name(fn, start)
    for _ in range(n_mutations):
        body = fn.body
        ret_at = max((i for i, st in enumerate(body) if isinstance(st, ast.Return)), default=len(body))
        first_def = next((i for i in range(start, len(body)) if name and any(
            isinstance(n, ast.Name) and n.id == name and isinstance(n.ctx, ast.Store) for n in ast.walk(body[i]))), None)
        tail_nodes = [n for st in body[start:] for n in ast.walk(st)]
        kind = rng.choice(["float", "float", "int", "insert", "insert", "drop", "swap", "op", "guard", "clamp", "cmp"])
        if kind == "float":
            cs = [n for n in tail_nodes if isinstance(n, ast.Constant) and type(n.value) is float and n.value != 0.0]
            for n in rng.sample(cs, min(len(cs), rng.randint(1, 3))):
                n.value = float(repr(round(n.value * rng.uniform(0.6, 1.6), 4))) or n.value
        elif kind == "int":
            cs = [n for n in tail_nodes if isinstance(n, ast.Constant) and type(n.value) is int and abs(n.value) >= 5]
            for n in rng.sample(cs, min(len(cs), rng.randint(1, 2))):
                n.value = max(5, int(n.value * rng.uniform(0.6, 1.6)))
        elif kind == "insert" and name and first_def is not None:
            sign = rng.choice(["+=", "+=", "-="])
            st = _parse_stmt(f"{name} {sign} {_fill(rng.choice(TERMS), rng)}")
            if rng.random() < 0.35:
                st = ast.If(test=_parse_expr(_fill(rng.choice(GUARDS), rng)), body=[st], orelse=[])
            body.insert(rng.randint(first_def + 1, ret_at), st)
        elif kind == "drop" and name:
            cand = [i for i in range(start, len(body)) if _is_update(body[i], name) or _only_updates(body[i], name)]
            if len(cand) > 1:
                del body[rng.choice(cand)]
        elif kind == "swap" and name:
            cand = [i for i in range(start, len(body) - 1) if (_is_update(body[i], name) or _only_updates(body[i], name))
                    and (_is_update(body[i + 1], name) or _only_updates(body[i + 1], name))]
            if cand:
                i = rng.choice(cand)
                body[i], body[i + 1] = body[i + 1], body[i]
        elif kind == "op":
            ops = [n for n in tail_nodes if isinstance(n, ast.BinOp) and (
                isinstance(n.op, (ast.Add, ast.Sub)) or
                (isinstance(n.op, ast.Mult) and isinstance(n.right, ast.Constant) and type(n.right.value) in (int, float)
                 and n.right.value != 0) or
                (isinstance(n.op, ast.Div) and isinstance(n.right, ast.Constant)))]
            if ops:
                n = rng.choice(ops)
                n.op = {ast.Add: ast.Sub, ast.Sub: ast.Add, ast.Mult: ast.Div, ast.Div: ast.Mult}[type(n.op)]()
        elif kind == "guard" and name:
            cand = [i for i in range(start, len(body)) if _is_update(body[i], name)]
            if cand:
                i = rng.choice(cand)
                body[i] = ast.If(test=_parse_expr(_fill(rng.choice(GUARDS), rng)), body=[body[i]], orelse=[])
        elif kind == "clamp" and name:
            ups = [n for n in tail_nodes if _is_update(n, name)]
            if ups:
                n = rng.choice(ups)
                how = rng.choice(["max", "min", "abs"])
                if how == "abs":
                    n.value = ast.Call(func=ast.Name(id="abs", ctx=ast.Load()), args=[n.value], keywords=[])
                else:
                    bound = _parse_expr(_coef(rng) if how == "min" else "0.0")
                    n.value = ast.Call(func=ast.Name(id=how, ctx=ast.Load()), args=[bound, n.value], keywords=[])
        elif kind == "cmp":
            cs = [n for n in tail_nodes if isinstance(n, ast.Compare) and len(n.ops) == 1
                  and isinstance(n.ops[0], (ast.Lt, ast.LtE, ast.Gt, ast.GtE))]
            if cs:
                n = rng.choice(cs)
                n.ops[0] = {ast.Lt: ast.LtE, ast.LtE: ast.Lt, ast.Gt: ast.GtE, ast.GtE: ast.Gt}[type(n.ops[0])]()
    _template_ending(fn)
    ast.fix_missing_locations(tree)
    return ast.unparse(tree) + "\n"


def policy_population(n: int, seed: int = 0, bases: Optional[Sequence[str]] = None, family_share: float = 0.08) -> List[str]:
    """n distinct policy sources, deterministic in `seed`: mutants (1-5 edits) of `bases` (default: the shipped sources
    and the template-style bodies), plus a share of the reference's random four-parameter family.  Every source passes
    the reference's validators; nothing is filtered by what the translator accepts."""
    rng = random.Random(seed)
    shipped = list(bases) if bases is not None else list(base_policies().values())
    out, seen = [], set()
    guard = 0
    while len(out) < n:
        guard += 1
        if guard > 50 * n + 100:
            raise RuntimeError("could not generate enough distinct policies")
        r = rng.random()
        if r < family_share:
            src = random_family_policy(rng)
        else:
            parent = template_body_policy(rng) if (bases is None and rng.random() < 0.3) else rng.choice(shipped)
            k = 1
            while k < 5 and rng.random() < 0.55:
                k += 1
            try:
                src = mutate_source(parent, rng, k)
            except (SyntaxError, ValueError, IndexError):
                continue
        if src in seen or not admitted_by_reference_validators(src):
            continue
        seen.add(src)
        out.append(src)
    return out
