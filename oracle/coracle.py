"""ctypes front end of the C oracle (TEST INFRASTRUCTURE; see ksim_oracle.c for what it restates).

`run(cluster, trace, policy, params)` takes the product's workload containers (workload.ClusterSpec /
workload.Trace) only as plain arrays; the built-in policies are the hand-translated ones."""
import ctypes as C

import numpy as np

from . import build_oracle

POLICY_IDS = {"first_fit": 0, "best_fit": 1, "funsearch_4901": 2, "funsearch_4816": 3, "funsearch_4800": 4,
              "worst_fit": 5, "random_family": 6}
# policy names of tests/golden/policies.json -> (built-in id, parameters)
GOLDEN_NAMES = {"ts_first_fit": "first_fit", "fs_first_fit": "first_fit", "ts_best_fit": "best_fit", "fs_best_fit": "best_fit",
                "ts_funsearch_4901": "funsearch_4901", "ts_funsearch_4816": "funsearch_4816",
                "ts_funsearch_4800": "funsearch_4800", "fs_worst_fit": "worst_fit"}


class OracleResult(C.Structure):
    _fields_ = [("status", C.c_int), ("placed", C.c_int), ("num_snapshots", C.c_int), ("num_frag", C.c_int),
                ("max_nodes", C.c_int), ("max_waiting", C.c_int), ("events", C.c_longlong), ("attempts", C.c_longlong),
                ("last_event_time", C.c_longlong), ("scorer_calls", C.c_longlong), ("avg_cpu", C.c_double),
                ("avg_mem", C.c_double), ("avg_gpu_count", C.c_double), ("avg_gpu_milli", C.c_double),
                ("fragmentation", C.c_double), ("score", C.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build_oracle.build())
        _lib.ks_oracle_run.restype = C.c_int
    return _lib


def id_ranks(pod_ids):
    order = sorted(range(len(pod_ids)), key=pod_ids.__getitem__)
    rank = np.empty(len(pod_ids), dtype=np.int32)
    rank[np.asarray(order, dtype=np.int64)] = np.arange(len(pod_ids), dtype=np.int32)
    return rank


def run(cluster, trace, policy: str, params=None, snapshot_interval: float = 0.05, max_events: int = 0):
    """Returns (OracleResult, final_time[int64], node_index[int32], gpu_mask[int32]) in pod_list order."""
    n, p = cluster.n_nodes, trace.n_pods
    a = lambda x, dt: np.ascontiguousarray(x, dtype=dt)  # noqa: E731
    cpu_t, mem_t, ng = a(cluster.cpu_total, np.int32), a(cluster.mem_total, np.int32), a(cluster.num_gpus, np.int32)
    gmt = a(cluster.gpu_milli_total, np.int32).reshape(n, 8)
    cpu, mem, ngpu, gm = (a(trace.cpu_milli, np.int32), a(trace.memory_mib, np.int32), a(trace.num_gpu, np.int32),
                          a(trace.gpu_milli, np.int32))
    ct, du = a(trace.creation_time, np.int64), a(trace.duration_time, np.int64)
    rank = id_ranks(list(trace.pod_ids))
    prm = a(params if params is not None else [0, 0, 0, 0], np.float64)
    res = OracleResult()
    o_t, o_n, o_m = np.zeros(max(p, 1), np.int64), np.zeros(max(p, 1), np.int32), np.zeros(max(p, 1), np.int32)
    vp = lambda x: x.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib().ks_oracle_run(C.c_int(n), vp(cpu_t), vp(mem_t), vp(ng), vp(gmt), C.c_longlong(p), vp(cpu), vp(mem), vp(ngpu),
                             vp(gm), vp(ct), vp(du), vp(rank), C.c_int(POLICY_IDS[policy]), vp(prm),
                             C.c_double(snapshot_interval), C.c_longlong(max_events), C.byref(res), vp(o_t), vp(o_n), vp(o_m))
    assert rc == 0
    return res, o_t[:p], o_n[:p], o_m[:p]


def digest(cluster, trace, final_time, node_index, gpu_mask) -> str:
    """Same byte format as pysim.placement_digest (placed pods ordered by (final time, id rank))."""
    import hashlib
    rank = id_ranks(list(trace.pod_ids))
    placed = np.nonzero(node_index >= 0)[0]
    order = placed[np.lexsort((rank[placed], final_time[placed]))]
    lines = []
    for pi in order:
        m = int(gpu_mask[pi])
        gl = "|".join(str(g) for g in range(8) if (m >> g) & 1)
        lines.append(f"{trace.pod_ids[pi]},{int(final_time[pi])},{cluster.node_ids[int(node_index[pi])]},{gl}")
    return hashlib.sha256("\n".join(lines).encode()).hexdigest()
