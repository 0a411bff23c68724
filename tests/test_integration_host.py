"""Host-side integration helpers on CPU: placement write-back and the invariant checker, fed with an
EpisodeOutcome built from the oracle's result (no GPU involved)."""
import numpy as np
import pytest

from helpers import load_golden, oracle_specs


def _outcome_from_oracle(res):
    from funsearch_kubernetes_simulator_b200.evaluator import EpisodeOutcome
    return EpisodeOutcome(res.status, res.score, None, res.events, res.attempts, res.scorer_calls, res.placed,
                          res.max_nodes, res.max_waiting, res.last_event_time, 0,
                          np.array(res.final_time, dtype=np.int64), np.array(res.node_index, dtype=np.int64),
                          np.array(res.gpu_mask, dtype=np.int64))


def test_write_back_and_invariants():
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.integration import check_invariants, write_back
    from oracle import pysim
    _g, pol = load_golden()
    cluster, trace = ClusterSpec.load("gpu16"), Trace.load("default").head(3000)
    cs, ts = oracle_specs(cluster, trace)
    res = pysim.run_episode(cs, ts, pysim.scorer_from_source(pol["ts_funsearch_4901"]))
    out = _outcome_from_oracle(res)
    check_invariants(out, cluster, trace)       # a feasible policy never over-commits

    class P:  # the reference's Pod fields that the simulator mutates
        pass
    pods = [P() for _ in range(trace.n_pods)]
    write_back(out, pods, cluster)
    for i in (0, 17, 2999):
        assert pods[i].creation_time == res.final_time[i]
        assert pods[i].assigned_node == cs.node_ids[res.node_index[i]]
        assert pods[i].assigned_gpus == sorted(res.gpu_lists[i])
    # placement order recovered from (final time, id rank)
    assert out.placement_order(trace.packed()[0]).tolist() == res.placement_order
    # a policy that ignores capacity over-commits a node: the checker must say so
    res2 = pysim.run_episode(cs, ts, pysim.scorer_from_source("def priority_function(pod, node):\n    return 5 if pod.num_gpu == 0 else 0\n"))
    with pytest.raises(ValueError):
        check_invariants(_outcome_from_oracle(res2), cluster, trace)
