"""The language the reference ADMITS, not the language this repository happens to translate: a corpus of policy sources
that pass the reference's two validators (funsearch/safe_execution.py:35-79, restated in variants.py) and together use
every construct of SURVEY.md App. B.  For each source the wrapped score int(max(0, f(pod, node))) of 256 random
(pod, node) cases must either equal CPython's exactly (raises included) or be DECLINED -- the translator refuses the source,
or the device reports 'unsupported' for the case.  A declined policy is one the reference would have scored and this
framework cannot (evaluate_population lists it under not_evaluated); what may never happen is a different number.
The count of declined sources is printed and capped; the cap tightens from round to round."""
import numpy as np
import pytest

from policy_fuzz import cpython_scores, random_cases

pytestmark = pytest.mark.gpu

DECLINED_CEILING = 0.10   # share of the corpus that may be declined as a whole (round 2: 4 of 50)

HEAD = "def priority_function(pod, node):\n"
GUARD = ("    if pod.cpu_milli > node.cpu_milli_left or pod.memory_mib > node.memory_mib_left or pod.num_gpu > node.gpu_left:\n"
         "        return 0\n")

CORPUS = {
    # --- statements ---
    "assign_augassign": HEAD + GUARD + "    score = 0.0\n    score += node.cpu_milli_left / node.cpu_milli_total * 100\n    score -= 3\n    score *= 1.5\n    score /= 2\n    return max(1, int(score))\n",
    "if_elif_else": HEAD + "    u = 1 - node.cpu_milli_left / node.cpu_milli_total\n    if u < 0.3:\n        s = 10\n    elif u < 0.7:\n        s = 20.5\n    else:\n        s = 5\n    return s\n",
    "for_over_gpus": HEAD + "    free = 0\n    for gpu in node.gpus:\n        if gpu.gpu_milli_left >= pod.gpu_milli:\n            free += 1\n    return free * 100 + 1\n",
    "for_range_enumerate": HEAD + "    s = 0\n    for i in range(3):\n        s += i * pod.num_gpu\n    for k, g in enumerate(node.gpus):\n        s += k * g.gpu_milli_left\n    return s + 1\n",
    "while_loop": HEAD + "    n = 0\n    left = node.cpu_milli_left\n    while left > pod.cpu_milli and n < 20 and pod.cpu_milli > 0:\n        left -= pod.cpu_milli\n        n += 1\n    return n + 1\n",
    "break_continue": HEAD + "    s = 0\n    for g in node.gpus:\n        if g.gpu_milli_left == 0:\n            continue\n        if g.gpu_milli_left < pod.gpu_milli:\n            break\n        s += g.gpu_milli_left\n    return s\n",
    "try_bare_except": HEAD + "    try:\n        r = node.cpu_milli_left / (node.gpu_left - pod.num_gpu)\n    except:\n        r = -1.0\n    return max(1, int(r + 10))\n",
    "try_except_else": HEAD + "    s = 5.0\n    try:\n        s += node.gpus[pod.num_gpu].gpu_milli_left\n    except:\n        s += 0.5\n    else:\n        s *= 2\n    return max(1, int(s))\n",
    "try_named_exception": HEAD + "    try:\n        r = pod.cpu_milli / node.gpu_left\n    except ZeroDivisionError:\n        r = 0.0\n    return max(1, int(r))\n",
    "nested_try": HEAD + "    s = 0.0\n    try:\n        try:\n            s = 1 / pod.num_gpu\n        except:\n            s = node.gpus[9].gpu_milli_left\n    except:\n        s = 7.25\n    return max(1, int(s * 4))\n",
    "assert_raise": HEAD + "    assert node.cpu_milli_total > 0\n    if pod.num_gpu > 4:\n        raise ValueError\n    return 3\n",
    "pass_docstring": HEAD + '    """docstring"""\n    pass\n    return 1\n',
    "tuple_unpack": HEAD + "    a, b = node.cpu_milli_left, node.memory_mib_left\n    a, b = b, a\n    return max(1, a - b)\n",
    "list_append_loop": HEAD + "    xs = []\n    for g in node.gpus:\n        if g.gpu_milli_left >= pod.gpu_milli:\n            xs.append(g.gpu_milli_left - pod.gpu_milli)\n    return (min(xs) if xs else 0) + len(xs) * 7 + 1\n",
    "round_two_digits": HEAD + "    u = round(node.cpu_milli_left / node.cpu_milli_total, 2)\n    v = round(node.memory_mib_left / 7, 3)\n    return max(1, int(u * 1000 + v))\n",
    "module_constant": "W_CPU = 0.7\nW_MEM = 0.3\n" + HEAD + "    return max(1, int((node.cpu_milli_left * W_CPU + node.memory_mib_left * W_MEM) / 10))\n",
    "early_return_none": HEAD + "    if pod.num_gpu > node.gpu_left:\n        return\n    return 5\n",
    # --- expressions ---
    "ifexp_boolop_not": HEAD + "    a = 10 if pod.num_gpu > 0 and not node.gpu_left == 0 else 2\n    b = pod.num_gpu or 7\n    c = node.gpu_left and 3\n    return a + b + c\n",
    "chained_compare": HEAD + "    u = node.memory_mib_left / node.memory_mib_total\n    return 50 if 0.2 < u <= 0.8 else 1\n",
    "all_binops": HEAD + "    a = node.cpu_milli_left\n    b = pod.cpu_milli + 1\n    return max(1, (a + b) - (a - b) + a * 2 // b + a % b + int(a / b) + (a // 7) ** 2)\n",
    "float_floor_mod": HEAD + "    x = node.cpu_milli_left / 7.0\n    return max(1, int(x // 3.5 + x % 2.25 + (-x) % 3))\n",
    "pow_variants": HEAD + "    u = node.cpu_milli_left / node.cpu_milli_total\n    return max(1, int(abs(u) ** 0.5 * 100 + 2 ** 5 + pod.num_gpu ** 2 + 10 ** -2 * 300 + (1 + abs(u)) ** 2.5))\n",
    "unary_ops": HEAD + "    return max(1, -(-node.gpu_left) + +pod.num_gpu + (~pod.num_gpu + 10))\n",
    "subscript_slice": HEAD + "    s = 0\n    if len(node.gpus) > 1:\n        s = node.gpus[0].gpu_milli_left + node.gpus[-1].gpu_milli_total\n    part = node.gpus[:pod.num_gpu]\n    return s + len(part) + len(node.gpus[1:])\n",
    "list_literal_index": HEAD + "    w = [0.5, 0.3, 0.2]\n    return max(1, int(node.cpu_milli_left * w[0] + node.memory_mib_left * w[1] + node.gpu_left * w[-1]))\n",
    "listcomp_filter": HEAD + "    fits = [g for g in node.gpus if g.gpu_milli_left >= pod.gpu_milli]\n    vals = [g.gpu_milli_left * 2 for g in fits]\n    return len(fits) * 10 + sum(vals) + 1\n",
    "genexp_reductions": HEAD + "    if not node.gpus:\n        return 1\n    return sum(g.gpu_milli_left for g in node.gpus) + max(g.gpu_milli_left for g in node.gpus) - min(g.gpu_milli_total for g in node.gpus) + 1000\n",
    "sorted_lambda_key": HEAD + "    t = sorted(node.gpus, key=lambda g: g.gpu_milli_left)\n    r = sorted([g.gpu_milli_left for g in node.gpus], reverse=True)\n    return (t[0].gpu_milli_left if t else 0) + (r[0] if r else 0) + 1\n",
    "builtin_calls": HEAD + "    x = node.cpu_milli_left / 3\n    return max(1, int(x)) + abs(-3) + len(node.gpus) + min(2, 3) + round(2.5) + round(x) + int(float(4)) + bool(x) + sum([1, 2])\n",
    "math_functions": HEAD + "    u = (node.cpu_milli_left + 1) / (node.cpu_milli_total + 1)\n    return max(1, int(math.sqrt(abs(u)) * 100 + math.log(1 + abs(u)) * 50 + math.exp(-abs(u)) * 10 + math.pow(1 + abs(u), 1.5)))\n",
    "operator_functions": HEAD + "    return max(1, int(operator.add(node.gpu_left, 2) * operator.truediv(node.cpu_milli_left, 1000) + operator.mod(node.memory_mib_left, 7) - operator.sub(3, 1) + operator.mul(2, 3)))\n",
    "max_min_first_extremal": HEAD + "    a = max(1, 1.0)\n    b = min(2.0, 2)\n    c = max(node.gpu_left, 0.0, pod.num_gpu)\n    return a + b + c\n",
    "sum_neumaier": HEAD + "    return max(1, int(sum([0.1, 0.2, 0.3, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1]) * 1e16 + sum(g.gpu_milli_left / 7 for g in node.gpus) * 1e12))\n",
    "int_float_compare": HEAD + "    big = node.memory_mib_total * 1000003\n    return 5 if big > 1e11 and node.gpu_left < 2.5 and 3 == 3.0 else 2\n",
    "all_fields": HEAD + "    s = pod.cpu_milli + pod.memory_mib + pod.num_gpu + pod.gpu_milli + pod.creation_time % 97 + pod.duration_time % 89\n    s += node.cpu_milli_left + node.memory_mib_left + node.gpu_left + node.cpu_milli_total + node.memory_mib_total\n    for g in node.gpus:\n        s += g.gpu_milli_left + g.gpu_milli_total + g.memory_mib_total + g.memory_mib_left\n    return max(1, s % 100000)\n",
    # --- things that raise in CPython (the worker scores the policy 0) ---
    "zero_division": HEAD + "    return 100 // (pod.num_gpu - pod.num_gpu)\n",
    "name_error_bare_sqrt": HEAD + "    return sqrt(4)\n",
    "index_error": HEAD + "    return node.gpus[8].gpu_milli_left\n",
    "type_error_none": HEAD + "    x = None\n    return x + 1\n",
    "int_of_nan_inf": HEAD + "    x = 1e308 * 10\n    return int(x - x)\n",
    "complex_from_negative_base": HEAD + "    return int((node.cpu_milli_left - node.cpu_milli_total - 1) ** 0.5)\n",
    "big_int": HEAD + "    return 10 ** 30 if node.gpu_left > 100 else (node.memory_mib_total * 1000) ** 2\n",
    # --- admitted by the validators, outside what is translated today (expected to be declined) ---
    "round_negative_digits": HEAD + "    return max(1, int(round(node.cpu_milli_left / 7, -2)))\n",
    "dict_literal": HEAD + "    w = {'cpu': 0.6, 'mem': 0.4}\n    return max(1, int(node.cpu_milli_left * w['cpu'] + node.memory_mib_left * w['mem']))\n",
    "set_membership": HEAD + "    return 5 if pod.num_gpu in {1, 2, 4} else 1\n",
    "fstring_str": HEAD + "    tag = f\"{pod.num_gpu}\"\n    return len(tag) + len(str(node.gpu_left))\n",
    "math_sin": HEAD + "    return max(1, int(100 + 50 * math.sin(node.cpu_milli_left / 1000)))\n",
    "list_extend_method": HEAD + "    xs = [1]\n    xs.extend([2, 3])\n    return sum(xs) + 1\n",
    "try_finally": HEAD + "    s = 1\n    try:\n        s = 10 // pod.num_gpu\n    except:\n        s = 2\n    finally:\n        s += 1\n    return s\n",
    "try_finally_propagates": HEAD + "    s = 1\n    try:\n        s = 10 // (pod.num_gpu - 1)\n    finally:\n        s += node.gpu_left\n    return s\n",
}


# FunSearchScheduler exec's whatever source it is given (funsearch_integration.py:77-89 never calls the validators), so
# hand-written policies may call their own helpers although validate_code_structure rejects calls to unknown names.
EXECUTED_NOT_ADMITTED = {
    "nested_def_helper": HEAD + "    def frac(a, b):\n        return a / max(1, b)\n    return max(1, int(frac(node.cpu_milli_left, node.cpu_milli_total) * 1000))\n",
    "lambda_helper": HEAD + "    half = lambda v: v / 2\n    return max(1, int(half(node.memory_mib_left)))\n",
    "module_level_helper": "K = 1.5\ndef util(left, total):\n    return (1 - left / total) * K\n" + HEAD +
                           "    return max(1, int(util(node.cpu_milli_left, node.cpu_milli_total) * 100 + util(node.memory_mib_left, node.memory_mib_total)))\n",
    "raise_with_call": HEAD + "    if pod.num_gpu > 4:\n        raise ValueError('too many')\n    return 3\n",
}


def test_corpus_is_admitted_by_the_reference_validators():
    from funsearch_kubernetes_simulator_b200.variants import admitted_by_reference_validators
    assert len(CORPUS) >= 40
    rejected = [k for k, src in CORPUS.items() if not admitted_by_reference_validators(src)]
    assert not rejected, rejected


def test_admitted_corpus_matches_cpython_or_is_declined(built_lib, capsys):
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    pods, ns, nc, objs = random_cases(256, seed=5)
    ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default").head(10)], processes=0)
    names = list(CORPUS)
    extra = list(EXECUTED_NOT_ADMITTED)
    compiled_extra = ev.compile([EXECUTED_NOT_ADMITTED[k] for k in extra], processes=0)
    assert all(c is not None for c in compiled_extra)
    for k, nm in enumerate(extra):
        got = ev.debug_scores(compiled_extra, k, pods, ns, nc)
        want = np.array(cpython_scores(EXECUTED_NOT_ADMITTED[nm], objs), dtype=np.int64)
        assert np.array_equal(got, want), nm
    compiled = ev.compile([CORPUS[k] for k in names], processes=0)
    declined_src, declined_cases, exact = [], {}, 0
    ok_idx = [i for i, c in enumerate(compiled) if c is not None]
    ok_pol = [compiled[i] for i in ok_idx]
    for k, i in enumerate(ok_idx):
        got = ev.debug_scores(ok_pol, k, pods, ns, nc)
        want = np.array(cpython_scores(CORPUS[names[i]], objs), dtype=np.int64)
        soft = got == -2
        bad = np.nonzero((got != want) & ~soft)[0]
        assert bad.size == 0, f"{names[i]}: case {int(bad[0])}: device {int(got[bad[0]])} vs CPython {int(want[bad[0]])}\n{CORPUS[names[i]]}"
        if soft.any():
            declined_cases[names[i]] = int(soft.sum())
        if not soft.all():
            exact += 1
    for i, c in enumerate(compiled):
        if c is None:
            declined_src.append((names[i], ev.compile_error(CORPUS[names[i]])))
    fully_declined = [n for n, k in declined_cases.items() if k == len(objs)]
    with capsys.disabled():
        print(f"\nadmitted corpus: {len(names)} sources; {exact} evaluated exactly; {len(declined_src)} declined by the translator "
              f"({', '.join(n for n, _ in declined_src)}); declined at run time for every case: {fully_declined}; "
              f"for some cases: { {n: k for n, k in declined_cases.items() if k < len(objs)} }")
    n_declined = len(declined_src) + len(fully_declined)
    assert n_declined <= DECLINED_CEILING * len(names), (n_declined, [n for n, _ in declined_src], fully_declined)
