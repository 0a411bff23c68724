"""Shared helpers for the test-suite: golden fixtures, oracle adapters, digest of a GPU outcome."""
import hashlib
import json
import os

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
GOLD = os.path.join(ROOT, "tests", "golden")


def load_golden():
    with open(os.path.join(GOLD, "golden_runs.json")) as f:
        g = json.load(f)
    with open(os.path.join(GOLD, "policies.json")) as f:
        p = json.load(f)
    return g, p


def oracle_specs(cluster, trace):
    """workload.ClusterSpec / workload.Trace -> the oracle's plain-list specs."""
    from oracle import pysim
    cs = pysim.ClusterSpec(list(cluster.node_ids), cluster.cpu_total.tolist(), cluster.mem_total.tolist(),
                           [[(int(cluster.gpu_mem_total[i, g]), int(cluster.gpu_milli_total[i, g]))
                             for g in range(int(cluster.num_gpus[i]))] for i in range(cluster.n_nodes)])
    ts = pysim.TraceSpec(list(trace.pod_ids), trace.cpu_milli.tolist(), trace.memory_mib.tolist(),
                         trace.num_gpu.tolist(), trace.gpu_milli.tolist(), trace.creation_time.tolist(),
                         trace.duration_time.tolist(), [""] * trace.n_pods)
    return cs, ts


def specs_from_edge(edge):
    from funsearch_kubernetes_simulator_b200.workload import ClusterSpec, Trace
    c, t = edge["cluster"], edge["trace"]
    n = len(c["node_ids"])
    gmil = np.zeros((n, 8), dtype=np.int32)
    gmem = np.zeros((n, 8), dtype=np.int32)
    for i, gl in enumerate(c["gpus"]):
        for j, (m, mt) in enumerate(gl):
            gmem[i, j], gmil[i, j] = m, mt
    cluster = ClusterSpec(c["node_ids"], c["cpu_total"], c["mem_total"], [len(g) for g in c["gpus"]], gmil, gmem)
    trace = Trace(t["pod_ids"], t["cpu_milli"], t["memory_mib"], t["num_gpu"], t["gpu_milli"], t["creation_time"],
                  t["duration_time"])
    return cluster, trace


def outcome_digest(trace, cluster, outcome):
    """Same byte format as oracle.pysim.placement_digest, from a GPU EpisodeOutcome."""
    rank = trace.packed()[0]
    lines = []
    for pi in outcome.placement_order(rank):
        m = int(outcome.gpu_mask[pi])
        gl = "|".join(str(g) for g in range(8) if (m >> g) & 1)
        lines.append(f"{trace.pod_ids[pi]},{int(outcome.final_creation_time[pi])},"
                     f"{cluster.node_ids[int(outcome.node_index[pi])]},{gl}")
    return hashlib.sha256("\n".join(lines).encode()).hexdigest()


def assert_matches_golden(tag, gold, outcome, trace, cluster, rel=1e-9):
    """Bit-exact integers and placement log; float metrics within 1e-9 relative (north-star tolerance)."""
    assert outcome.status == gold["status"], f"{tag}: status {outcome.status_name} vs golden {gold['status']} ({gold['error']})"
    assert outcome.events == gold["events"], f"{tag}: events {outcome.events} vs {gold['events']}"
    assert outcome.attempts == gold["attempts"], f"{tag}: attempts"
    if gold["status"] != 0:
        assert outcome.policy_score == 0.0
        return
    assert outcome.placed == gold["placed"], f"{tag}: placed"
    assert outcome.scorer_calls == gold["scorer_calls"], f"{tag}: scorer_calls"
    assert outcome.max_nodes == gold["max_nodes"], f"{tag}: max_nodes {outcome.max_nodes} vs {gold['max_nodes']}"
    assert outcome.max_waiting == gold["max_waiting"], f"{tag}: max_waiting"
    assert outcome.last_event_time == gold["last_event_time"], f"{tag}: last_event_time"
    ns = outcome.results.num_snapshots if outcome.results else 0
    nf = outcome.results.num_fragmentation_events if outcome.results else 0
    assert ns == gold["num_snapshots"], f"{tag}: snapshots {ns} vs {gold['num_snapshots']}"
    if outcome.results:
        assert nf == gold["num_fragmentation_events"], f"{tag}: fragmentation events"
    if outcome.final_creation_time is not None:
        assert outcome_digest(trace, cluster, outcome) == gold["digest"], f"{tag}: placement log differs"
    if gold["has_results"]:
        r = outcome.results
        for name, got in (("avg_cpu", r.avg_cpu_utilization), ("avg_mem", r.avg_memory_utilization),
                          ("avg_gpu_count", r.avg_gpu_count_utilization), ("avg_gpu_milli", r.avg_gpu_memory_utilization),
                          ("fragmentation", r.gpu_fragmentation_score)):
            want = float.fromhex(gold[name])
            assert abs(got - want) <= rel * max(abs(want), 1e-300), f"{tag}: {name} {got!r} vs {want!r}"
    want = float.fromhex(gold["score"])
    assert abs(outcome.policy_score - want) <= rel * max(abs(want), 1e-300), f"{tag}: score {outcome.policy_score!r} vs {want!r}"
