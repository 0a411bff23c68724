"""The C oracle (oracle/ksim_oracle.c) against the golden fixtures recorded from the reference: it must be
bit-identical on placement logs and counters before it is trusted at sizes the Python oracle cannot reach."""
import pytest

from helpers import load_golden


@pytest.fixture(scope="module")
def gold():
    return load_golden()


def _check(tag, g, res, cluster, trace, t, n, m):
    from oracle import coracle
    assert res.status == g["status"], tag
    assert (res.events, res.attempts, res.placed) == (g["events"], g["attempts"], g["placed"]), tag
    assert res.scorer_calls == g["scorer_calls"] and res.max_nodes == g["max_nodes"] and res.max_waiting == g["max_waiting"], tag
    assert res.num_snapshots == g["num_snapshots"] and res.num_frag == g["num_fragmentation_events"], tag
    assert res.last_event_time == g["last_event_time"], tag
    assert coracle.digest(cluster, trace, t, n, m) == g["digest"], tag
    for k, v in (("avg_cpu", res.avg_cpu), ("avg_mem", res.avg_mem), ("avg_gpu_count", res.avg_gpu_count),
                 ("avg_gpu_milli", res.avg_gpu_milli), ("fragmentation", res.fragmentation), ("score", res.score)):
        want = float.fromhex(g[k])
        assert abs(v - want) <= 1e-12 * max(abs(want), 1e-300), (tag, k, v, want)


def test_c_oracle_reproduces_reference_goldens(gold):
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from oracle import coracle
    g, _pol = gold
    cluster = ClusterSpec.load("gpu16")
    for tname in ("default", "cpu050", "cpu300", "gpushare40", "gpuspec25"):
        trace = Trace.load(tname)
        for gname, builtin in coracle.GOLDEN_NAMES.items():
            tag = f"{tname}|gpu16|{gname}"
            res, t, n, m = coracle.run(cluster, trace, builtin)
            _check(tag, g["runs"][tag], res, cluster, trace, t, n, m)


def test_c_oracle_random_family_and_other_clusters(gold):
    import re
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from oracle import coracle
    g, pol = gold
    cluster = ClusterSpec.load("gpu16")
    trace = Trace.load("default")
    for name in ("fs_random0", "fs_random1"):
        src = pol[name]
        base, cpu_f, mem_f = [float(x) for x in re.search(r"score = (\S+) \+ node.cpu_milli_left \* (\S+) \+ node.memory_mib_left \* (\S+)", src).groups()]
        gpu_f = float(re.search(r"score \+= node.gpu_left \* (\S+)", src).group(1))
        res, t, n, m = coracle.run(cluster, trace, "random_family", [cpu_f, mem_f, gpu_f, base])
        _check(name, g["runs"][f"default|gpu16|{name}"], res, cluster, trace, t, n, m)
    big = ClusterSpec.load("all1523").head(1024)
    for gname in ("ts_first_fit", "ts_best_fit", "ts_funsearch_4816"):
        res, t, n, m = coracle.run(big, trace, coracle.GOLDEN_NAMES[gname])
        _check(gname, g["runs"][f"default|all1024|{gname}"], res, big, trace, t, n, m)
    c64 = ClusterSpec.load("gpu1213").head(64)
    tr2 = Trace.load("cpu300").head(4000)
    for gname in ("ts_best_fit", "ts_funsearch_4901", "ts_funsearch_4800"):
        res, t, n, m = coracle.run(c64, tr2, coracle.GOLDEN_NAMES[gname])
        _check(gname, g["runs"][f"cpu300[:4000]|gpu64|{gname}"], res, c64, tr2, t, n, m)
