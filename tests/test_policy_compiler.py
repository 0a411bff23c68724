"""CPU tests of the Python -> CUDA C policy translator: every shipped policy translates, NVRTC accepts
the output for sm_100a, everything links against the replay kernel; unsupported constructs are refused."""
import ctypes as C

import pytest

from helpers import load_golden


def test_shipped_policies_translate_compile_and_link(built_lib):
    from funsearch_kubernetes_simulator_b200 import _native as nat
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    _g, pol = load_golden()
    every = [compile_policy(src) for src in pol.values()]
    cps = list({c.symbol: c for c in every}.values())   # one device function per STRUCTURE
    assert 10 <= len(cps) < len(every)
    n = len(cps)
    srcs = (C.c_char_p * n)(*[c.cuda_source.encode() for c in cps])
    syms = (C.c_char_p * n)(*[c.symbol.encode() for c in cps])
    flags = (C.c_longlong * n)(*[c.flags for c in cps])
    handles = (C.c_longlong * n)()
    status = (C.c_int * n)()
    nat.check(nat.lib().ks_policy_compile_many(n, srcs, syms, flags, 0, handles, status))
    assert list(status) == [0] * n, nat.lib().ks_compile_log().decode()
    size = C.c_longlong(0)
    nat.check(nat.lib().ks_program_link_check(n, handles, C.byref(size)))
    assert size.value > 100000
    # cache: same source -> same handle, no recompilation
    again = C.c_longlong(-1)
    nat.check(nat.lib().ks_policy_compile(cps[0].cuda_source.encode(), cps[0].symbol.encode(), cps[0].flags, C.byref(again)))
    assert again.value == handles[0]


def test_policies_that_differ_in_float_literals_share_a_structure():
    """The float literals are lifted into the per-policy constant vector (policy ABI v3): sources that differ only in
    them translate to the same CUDA text and symbol; int literals, operators and names are structure."""
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    _g, pol = load_golden()
    a, b = compile_policy(pol["fs_random0"]), compile_policy(pol["fs_random1"])
    assert a.symbol == b.symbol and a.cuda_source == b.cuda_source and a.consts != b.consts and len(a.consts) == 4
    base = "def priority_function(pod, node):\n    w = [0.25, 2.5]\n    s = 0.0\n    s += node.cpu_milli_left * 1.5 + w[1]\n    return max(1, int(s * {k}))\n"
    x, y, z = (compile_policy(base.format(k=k)) for k in ("3.0", "7.25", "3"))
    assert x.symbol == y.symbol and x.consts == (0.25, 2.5, 1.5, 3.0) and y.consts == (0.25, 2.5, 1.5, 7.25)
    assert z.symbol == x.symbol and z.consts == x.consts    # an int literal used as a float (s * 3) is lifted as 3.0
    i0, i1 = (compile_policy(base.format(k="3.0").replace("s * 3.0", f"s * 3.0 + pod.num_gpu * {m}")) for m in (4, 5))
    assert i0.symbol != i1.symbol and i0.consts == i1.consts                # an int that stays an int is structure
    assert "0x" not in x.cuda_source.split("{", 1)[1].replace("0x7ff", "")  # no float literal left in the code but 0.0
    assert compile_policy(base.format(k="3.0").replace("+ w[1]", "- w[1]")).symbol != x.symbol
    # module-level constants are lifted too, after the function's own literals
    m = compile_policy("ALPHA = 0.75\nBETA = [2, 0.5]\ndef priority_function(pod, node):\n    return node.gpu_left * ALPHA + BETA[1] + 1.25 + BETA[0]\n")
    assert m.consts == (1.25, 0.75, 0.5)
    # values that only exist as bit patterns survive the trip
    e = compile_policy("def priority_function(pod, node):\n    return 1e999 if node.gpu_left > 5e-324 else 1.7976931348623157e308\n")
    assert e.consts == (5e-324, float("inf"), 1.7976931348623157e308)   # AST field order: test, body, orelse


def test_nvrtc_error_is_reported(built_lib):
    from funsearch_kubernetes_simulator_b200 import _native as nat
    h = C.c_longlong(0)
    rc = nat.lib().ks_policy_compile(b'extern "C" __device__ long long bad(KS_POLICY_PARAMS) { return nope; }', b"bad", 0, C.byref(h))
    assert rc != 0 and h.value == -1
    assert b"nope" in nat.lib().ks_compile_log()


def test_static_types_of_the_champions():
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    _g, pol = load_golden()
    t = compile_policy(pol["ts_funsearch_4901"]).var_types
    assert t["score"] == "float" and t["gpu_score"] == "num" and t["free_gpu_millis"] == "int"
    t = compile_policy(pol["ts_funsearch_4800"]).var_types
    assert t["viable_gpus"] == "gpulist" and t["gpu_eff"] == "float"
    src = compile_policy(pol["ts_funsearch_4816"]).cuda_source
    assert "ks_fpow" in src and "ks_div" in src and "fma" not in src.lower()


def test_template_prefix_is_hoisted_only_when_it_is_exactly_the_template():
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    _g, pol = load_golden()
    assert compile_policy(pol["ts_funsearch_4901"]).flags == 1   # re-binds `gpu` in generator expressions only
    assert compile_policy(pol["fs_first_fit"]).flags == 1
    assert compile_policy(pol["x_ignores_gpu_milli"]).flags == 0
    src = pol["fs_first_fit"]
    assert compile_policy(src.replace("return 1000", "return 1000 + available_gpus")).flags == 0  # reads a prefix variable
    assert compile_policy(src.replace("pod.cpu_milli > node.cpu_milli_left", "pod.cpu_milli >= node.cpu_milli_left")).flags == 0
    renamed = src.replace("available_gpus", "n_ok").replace("gpu", "g").replace("g_milli", "gpu_milli").replace("num_g", "num_gpu").replace("node.gs", "node.gpus").replace("g_left", "gpu_left")
    assert compile_policy(renamed).flags == 1


@pytest.mark.parametrize("body", [
    "import os\n    return 1",
    "node.cpu_milli_left = 0\n    return 1",
    "try:\n        return 1\n    finally:\n        pass",
    "return round(pod.cpu_milli / 3, -2)",
    "return len(str(pod.cpu_milli))",
    "def helper(x):\n        y = x\n        return y\n    return helper(1)",
    "d = {}\n    d['a'] = 1\n    return 1",
    "w = {'a': 1}\n    return sum(w.values())",
    "return sum(x for x in {1, 2})",
])
def test_unsupported_constructs_are_refused(body):
    from funsearch_kubernetes_simulator_b200.policy_compiler import PolicyCompileError, compile_policy
    with pytest.raises(PolicyCompileError):
        compile_policy("def priority_function(pod, node):\n    " + body + "\n")


def test_widened_constructs_translate():
    """try / bare except / else, single-return helper functions (nested and module level), round(x, n), lists built with
    .append(): translated, not refused."""
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    t = compile_policy("def priority_function(pod, node):\n    try:\n        s = 1 / pod.num_gpu\n    except:\n        s = 2.5\n"
                       "    else:\n        s += 1\n    return s\n").cuda_source
    assert "goto ks_try1" in t and "st = 0;" in t
    # a named exception class does not exist in the sandbox: the clause raises NameError while matching, nothing is caught
    t = compile_policy("def priority_function(pod, node):\n    try:\n        s = 1 / pod.num_gpu\n    except ZeroDivisionError:\n        s = 2.5\n    return s\n").cuda_source
    assert "goto" not in t
    t = compile_policy("W = 3\ndef f(a):\n    return a * W\ndef priority_function(pod, node):\n    def g(b):\n        return b + 1\n    return g(f(pod.num_gpu))\n").cuda_source
    assert "3LL" in t
    assert "ks_fround_n" in compile_policy("def priority_function(pod, node):\n    return round(node.gpu_left / 3, 2)\n").cuda_source
    cp = compile_policy("def priority_function(pod, node):\n    xs = []\n    for g in node.gpus:\n        xs.append(g.gpu_milli_left * 0.5)\n    return sum(xs)\n")
    assert cp.var_types["xs"] == "flist"


def test_runtime_errors_become_exception_returns():
    """NameError / ZeroDivisionError / TypeError sites translate to KS_RET_EXCEPTION, not to compile errors."""
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    for body in ("return sqrt(4)", "return 1 / 0", "return max(1, None)", "return node.gpus[0].gpu_milli_left",
                 "x = [1, 2]\n    return x[5]", "return math.floor(2.5)", "return pod.nonexistent"):
        src = compile_policy("def priority_function(pod, node):\n    " + body + "\n").cuda_source
        assert "KS_RET_EXCEPTION" in src, body


def test_variant_population_is_deterministic_and_translatable():
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    a = policy_population(64, seed=3)
    assert a == policy_population(64, seed=3) and len(set(a)) == 64
    assert a != policy_population(64, seed=4)


def test_compile_pool_translates_compiles_and_links_without_a_gpu(built_lib, tmp_path):
    """The two phases a fresh generation goes through in the worker processes: sources -> (structure, float literals),
    then new structures -> NVRTC programs -> one linked, loadable module image per worker.  No GPU is needed for any of it."""
    from funsearch_kubernetes_simulator_b200 import compile_pool
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    pop = policy_population(24, seed=9) + ["def priority_function(pod, node):\n    return len(f'{pod.num_gpu}')\n"]
    try:
        compile_pool.start(2, 8)
        info = compile_pool.wait_ready()
        assert info["workers"] == 2
        res, secs = compile_pool.run_jobs("translate", [pop[:12], pop[12:]])
        assert all(r is not None for r in res) and len(secs) == 2
        flat = res[0] + res[1]
        assert flat[-1][0] == "error" and "JoinedStr" in flat[-1][1]
        for src, r in zip(pop[:-1], flat[:-1]):
            cp = compile_policy(src)
            assert (r[1], r[2], tuple(r[3])) == (cp.symbol, cp.flags, cp.consts) and r[0] == cp.cuda_source
        structs = list({r[1]: (r[0], r[1], r[2]) for r in flat[:-1]}.values())
        groups = [structs[i:i + 8] for i in range(0, len(structs), 8)]
        img = str(tmp_path / "module.cubin")
        out, _ = compile_pool.run_jobs("nvrtc_link", [(groups, img)])
        assert out[0]["linked"] and all(x == 0 for st in out[0]["status"] for x in st)
        import os
        assert os.path.getsize(img) > 100000 and open(img, "rb").read(4) == b"\x7fELF"
    finally:
        compile_pool.shutdown()
