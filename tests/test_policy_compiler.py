"""CPU tests of the Python -> CUDA C policy translator: every shipped policy translates, NVRTC accepts
the output for sm_100a, everything links against the replay kernel; unsupported constructs are refused."""
import ctypes as C

import pytest

from helpers import load_golden


def test_shipped_policies_translate_compile_and_link(built_lib):
    from funsearch_kubernetes_simulator_b200 import _native as nat
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    _g, pol = load_golden()
    cps = [compile_policy(src) for src in pol.values()]
    assert len({c.symbol for c in cps}) == len(set(pol.values()))
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
    "try:\n        return 1\n    except:\n        return 2",
    "return round(pod.cpu_milli / 3, 2)",
    "return len(str(pod.cpu_milli))",
    "def helper(x):\n        return x\n    return helper(1)",
    "d = {}\n    return 1",
])
def test_unsupported_constructs_are_refused(body):
    from funsearch_kubernetes_simulator_b200.policy_compiler import PolicyCompileError, compile_policy
    with pytest.raises(PolicyCompileError):
        compile_policy("def priority_function(pod, node):\n    " + body + "\n")


def test_runtime_errors_become_exception_returns():
    """NameError / ZeroDivisionError / TypeError sites translate to KS_RET_EXCEPTION, not to compile errors."""
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    for body in ("return sqrt(4)", "return 1 / 0", "return max(1, None)", "return node.gpus[0].gpu_milli_left",
                 "x = [1, 2]\n    return x[5]", "return math.floor(2.5)", "return pod.nonexistent"):
        src = compile_policy("def priority_function(pod, node):\n    " + body + "\n").cuda_source
        assert "KS_RET_EXCEPTION" in src, body


def test_variant_population_is_deterministic_and_translatable():
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    a = policy_population(64, seed=3)
    assert a == policy_population(64, seed=3) and len(set(a)) == 64
    assert a != policy_population(64, seed=4)
