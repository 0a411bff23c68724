"""ctypes binding of libks_b200.so (the C ABI declared in include/ks_b200.h).

There is deliberately no fallback: if the shared library is missing, or a GPU entry point is used
without a B200, the call raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# KS_B200_LIB selects another build of the library (e.g. libks_b200_invariants.so, the debug variant of build.py)
LIB_PATH = os.environ.get("KS_B200_LIB") or os.path.join(_HERE, "libks_b200.so")

# status words (include/ks_abi.h)
ST_OK, ST_POLICY_EXCEPTION, ST_GPU_SHORTFALL, ST_STEP_BUDGET, ST_UNSUPPORTED, ST_CAPACITY, ST_INVARIANT = range(7)
STATUS_NAMES = {0: "ok", 1: "policy_exception", 2: "gpu_shortfall", 3: "step_budget", 4: "unsupported", 5: "capacity",
                6: "invariant_violated"}
BUILD_PHASE_CLOCKS, BUILD_INVARIANTS = 1, 2

KS_POD_INTS = 8
KS_NC_FIELDS = 20
KS_NS_FIELDS = 12
KS_NODE_NONE = 0xFFFF
KS_NODE_WAITING = 0xFFFE
KS_SCRATCH_HEADER = 4096
KS_MAX_TIME = (1 << 39) - 2
KS_STATIC_SCRATCH_BYTES = 32 << 20
KS_PF_PREFIX_HOISTED = 1
KS_PF_READS_TIME = 2
KS_PF_READS_DURATION = 4
KS_MAX_TABLE = 4096
KS_API_VERSION = 3

# struct KsResult, 128 bytes
RESULT_DTYPE = np.dtype([
    ("status", np.int32), ("placed", np.int32), ("num_snapshots", np.int32),
    ("num_fragmentation_events", np.int32), ("max_nodes", np.int32), ("max_waiting", np.int32),
    ("events", np.int64), ("attempts", np.int64), ("last_event_time", np.int64),
    ("avg_cpu_utilization", np.float64), ("avg_memory_utilization", np.float64),
    ("avg_gpu_count_utilization", np.float64), ("avg_gpu_memory_utilization", np.float64),
    ("gpu_fragmentation_score", np.float64), ("policy_score", np.float64),
    ("scorer_calls", np.int64), ("device_cycles", np.int64), ("reserved", np.int64, (2,)),
])
assert RESULT_DTYPE.itemsize == 128


class KsPlan(C.Structure):
    _fields_ = [("n_episodes", C.c_int), ("n_nodes", C.c_int), ("max_pods", C.c_int), ("want_placements", C.c_int),
                ("warps_per_team", C.c_int), ("teams_per_block", C.c_int), ("heap_smem_entries", C.c_int), ("wait_smem_entries", C.c_int),
                ("ctas_per_sm", C.c_int), ("blocks", C.c_int), ("smem_per_team", C.c_int), ("smem_per_block", C.c_int),
                ("scratch_per_team", C.c_longlong), ("scratch_bytes", C.c_longlong)]


class KsBatch(C.Structure):
    _fields_ = [("traces", C.c_void_p), ("ep_policy", C.c_void_p), ("ep_const", C.c_void_p), ("consts", C.c_void_p),
                ("ep_trace", C.c_void_p), ("range_begin", C.c_void_p), ("nconst", C.c_void_p),
                ("results", C.c_void_p), ("placements", C.c_void_p), ("place_off", C.c_void_p), ("scratch", C.c_void_p),
                ("totals", C.c_longlong * 4), ("max_events", C.c_longlong), ("snapshot_interval", C.c_double)]


class KsHostTrace(C.Structure):
    _fields_ = [("podrec", C.c_void_p), ("heap0", C.c_void_p), ("place0", C.c_void_p), ("n_pods", C.c_longlong)]


class KsHostBatch(C.Structure):
    _fields_ = [("traces", C.POINTER(KsHostTrace)), ("n_traces", C.c_int), ("reserved", C.c_int),
                ("ep_policy", C.c_void_p), ("ep_const", C.c_void_p), ("consts", C.c_void_p), ("n_consts", C.c_longlong),
                ("ep_trace", C.c_void_p), ("range_begin", C.c_void_p), ("nconst", C.c_void_p), ("results", C.c_void_p),
                ("placements", C.c_void_p), ("place_off", C.c_void_p), ("placements_len", C.c_longlong),
                ("totals", C.c_longlong * 4), ("max_events", C.c_longlong), ("snapshot_interval", C.c_double)]


class KsError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libks_b200 error {code}: {message}")
        self.code = code


_lib = None

_SIGNATURES = {
    "ks_last_error": (C.c_char_p, []),
    "ks_compile_log": (C.c_char_p, []),
    "ks_api_version": (C.c_int, []),
    "ks_build_flags": (C.c_int, []),
    "ks_init": (C.c_int, [C.c_int]),
    "ks_trace_pack": (C.c_int, [C.c_longlong] + [C.c_void_p] * 10),
    "ks_cluster_pack": (C.c_int, [C.c_int] + [C.c_void_p] * 7),
    "ks_policy_compile": (C.c_int, [C.c_char_p, C.c_char_p, C.c_longlong, C.POINTER(C.c_longlong)]),
    "ks_policy_compile_many": (C.c_int, [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), C.POINTER(C.c_longlong),
                                         C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_int)]),
    "ks_set_cache_dir": (C.c_int, [C.c_char_p]),
    "ks_set_compile_batch": (C.c_int, [C.c_int]),
    "ks_program_link": (C.c_int, [C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "ks_program_link_check": (C.c_int, [C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "ks_program_release": (C.c_int, [C.c_longlong]),
    "ks_program_link_image": (C.c_int, [C.c_int, C.POINTER(C.c_longlong), C.c_char_p]),
    "ks_program_load_image": (C.c_int, [C.c_void_p, C.c_longlong, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "ks_plan": (C.c_int, [C.POINTER(KsPlan)]),
    "ks_eval_batch": (C.c_int, [C.c_longlong, C.POINTER(KsPlan), C.POINTER(KsBatch), C.c_void_p]),
    "ks_eval_batch_static": (C.c_int, [C.c_longlong, C.POINTER(KsPlan), C.POINTER(KsBatch), C.c_void_p]),
    "ks_host_workspace_bytes": (C.c_int, [C.POINTER(KsPlan), C.POINTER(KsHostBatch), C.POINTER(C.c_longlong)]),
    "ks_eval_batch_host": (C.c_int, [C.c_longlong, C.POINTER(KsPlan), C.POINTER(KsHostBatch), C.c_void_p, C.c_longlong,
                                     C.c_void_p, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "ks_debug_pow": (C.c_int, [C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "ks_debug_math": (C.c_int, [C.c_longlong, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "ks_debug_score": (C.c_int, [C.c_longlong, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p]),
    "ks_launch_count": (C.c_longlong, []),
    "ks_shutdown": (C.c_int, []),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def lib() -> C.CDLL:
    """The loaded library; raises if it has not been built (python __graft_entry__.py / build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise KsError(-1, f"{LIB_PATH} is missing: build it with funsearch_kubernetes_simulator_b200/build.py "
                              "(there is no Python/CPU fallback for the replay kernel)")
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype = res
            fn.argtypes = args
        _lib = h
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise KsError(rc, lib().ks_last_error().decode(errors="replace"))


def ptr(a: np.ndarray) -> int:
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data
