"""CPU tests (-m "not gpu"): the oracle against the golden fixtures recorded from the reference, the host
packing logic, the policy translator, and the C ABI surface (no compute calls: there is no GPU here)."""
import ctypes as C
import heapq
import json
import os
import re

import numpy as np
import pytest

from helpers import ROOT, load_golden, oracle_specs, specs_from_edge


@pytest.fixture(scope="module")
def gold():
    return load_golden()


def _check_oracle(tag, g, res, cs, ts):
    from oracle import pysim
    assert res.status == g["status"], tag
    assert res.events == g["events"] and res.attempts == g["attempts"], tag
    if g["status"] != 0:
        return
    assert res.placed == g["placed"] and res.scorer_calls == g["scorer_calls"], tag
    assert res.num_snapshots == g["num_snapshots"] and res.num_fragmentation_events == g["num_fragmentation_events"], tag
    assert res.max_nodes == g["max_nodes"] and res.max_waiting == g["max_waiting"], tag
    assert pysim.placement_digest(ts, cs, res) == g["digest"], tag
    assert pysim.placement_digest_allocation_order(ts, cs, res) == g["digest_alloc_order"], tag
    assert res.score.hex() == g["score"], tag
    for k, v in (("avg_cpu", res.avg_cpu), ("avg_mem", res.avg_mem), ("avg_gpu_count", res.avg_gpu_count),
                 ("avg_gpu_milli", res.avg_gpu_milli), ("fragmentation", res.fragmentation)):
        assert v.hex() == g[k], (tag, k)


def test_oracle_reproduces_survey_appendix_d(gold):
    """The five tests/test_scheduler.py policies on the default trace: placement-log SHA-256 prefixes and
    scores as recorded from the reference in SURVEY.md App. D (README.md:27-29 to printed precision)."""
    g, pol = gold
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from oracle import pysim
    cs, ts = oracle_specs(ClusterSpec.load("gpu16"), Trace.load("default"))
    want = {"ts_first_fit": ("c4966e1aead5abda", 19456, 0.42920557012850735),
            "ts_best_fit": ("531fe78725838271", 16383, 0.44654782731316534),
            "ts_funsearch_4901": ("55d23300731645cb", 27563, 0.49013357497851473),
            "ts_funsearch_4816": ("5c34c05343194658", 18657, 0.48162002959239419),
            "ts_funsearch_4800": ("ea40b3d911871b7d", 18603, 0.48002299859874331)}
    for name, (sha, events, score) in want.items():
        res = pysim.run_episode(cs, ts, pysim.scorer_from_source(pol[name]))
        assert pysim.placement_digest_allocation_order(ts, cs, res)[:16] == sha
        assert res.events == events and res.placed == 8152 and res.max_nodes == 16
        assert res.score == score
        _check_oracle(name, g["runs"][f"default|gpu16|{name}"], res, cs, ts)


def test_oracle_on_other_traces_and_raising_policies(gold):
    g, pol = gold
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from oracle import pysim
    cluster = ClusterSpec.load("gpu16")
    for tname, pname in (("cpu050", "ts_funsearch_4800"), ("cpu300", "ts_best_fit"), ("gpushare40", "fs_gpu_aware"),
                         ("default", "x_ignores_capacity"), ("default", "x_div_by_len_gpus")):
        cs, ts = oracle_specs(cluster, Trace.load(tname))
        res = pysim.run_episode(cs, ts, pysim.scorer_from_source(pol[pname]))
        _check_oracle(f"{tname}|{pname}", g["runs"][f"{tname}|gpu16|{pname}"], res, cs, ts)


def test_oracle_edge_cases(gold):
    g, pol = gold
    from oracle import pysim
    for name, edge in g["edge_cases"].items():
        cluster, trace = specs_from_edge(edge)
        cs, ts = oracle_specs(cluster, trace)
        res = pysim.run_episode(cs, ts, pysim.scorer_from_source(pol[edge["policy"]]))
        _check_oracle(name, g["runs"][f"edge:{name}|{edge['policy']}"], res, cs, ts)
    # the reference's tests/test_simulator.py known answer
    edge = g["edge_cases"]["test_simulator"]
    cluster, trace = specs_from_edge(edge)
    cs, ts = oracle_specs(cluster, trace)
    res = pysim.run_episode(cs, ts, pysim.scorer_from_source(pol["sim_int_best_fit"]))
    assert [cs.node_ids[i] for i in res.node_index] == ["node2", "node1", "node2", "node1"]
    assert res.gpu_lists == [[], [0], [], [0, 1]]
    assert res.cpu_left == cs.cpu_total and res.mem_left == cs.mem_total


# ------------------------------------------------------------------------------------------ C ABI surface
def test_library_exports_every_declared_symbol(built_lib):
    hdr = open(os.path.join(ROOT, "include", "ks_b200.h")).read()
    declared = set(re.findall(r"^(?:const char\*|int|long long)\s+(ks_\w+)\(", hdr, flags=re.M))
    assert len(declared) >= 18
    lib = C.CDLL(built_lib)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/ks_b200.h but not exported"
    from funsearch_kubernetes_simulator_b200 import _native as nat
    assert declared == set(nat.EXPORTED_SYMBOLS)
    assert nat.lib().ks_api_version() == nat.KS_API_VERSION == 3


def test_no_gpu_means_loud_failure(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, _native as nat
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    with pytest.raises(nat.KsError):
        BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
    assert nat.lib().ks_init(0) != 0
    assert b"no CPU fallback" in nat.lib().ks_last_error()


def test_trace_pack_matches_heapq_heapify(built_lib):
    """ks_trace_pack's initial heap image == CPython heapq.heapify over (time, pod_id) (event_simulator.py:23-34),
    incl. an unsorted trace with mixed-width ids (string order != numeric order)."""
    from funsearch_kubernetes_simulator_b200 import Trace
    rng = np.random.default_rng(5)
    for name in ("default", "cpu300", "cpu050"):
        tr = Trace.load(name)
        rank, podrec, heap0, place0 = tr.packed()
        items = [(int(t), pid, i) for i, (t, pid) in enumerate(zip(tr.creation_time, tr.pod_ids))]
        heapq.heapify(items)
        want = [(t << 24) | (int(rank[i]) << 1) for (t, _pid, i) in items]
        assert heap0[:tr.n_pods].tolist() == want, name
        # rank = string order
        order = sorted(range(tr.n_pods), key=tr.pod_ids.__getitem__)
        assert [int(rank[i]) for i in order] == list(range(tr.n_pods))
        # pod records are indexed by rank
        i = int(rng.integers(tr.n_pods))
        rec = podrec[int(rank[i]) * 8: int(rank[i]) * 8 + 8]
        assert rec[0] == tr.cpu_milli[i] and rec[1] == tr.memory_mib[i] and rec[2] == tr.num_gpu[i] and rec[3] == tr.gpu_milli[i]
        assert (int(rec[4]) & 0xFFFFFFFF) | (int(rec[5]) << 32) == tr.duration_time[i] and rec[6] == i
        assert int(place0[int(rank[i])]) == (int(tr.creation_time[i]) << 24) | (0xFFFF << 8)


def test_trace_pack_rejects_bad_input(built_lib):
    from funsearch_kubernetes_simulator_b200 import Trace, _native as nat
    with pytest.raises(ValueError):
        Trace(["a", "a"], [1, 1], [1, 1], [0, 0], [0, 0], [0, 1], [1, 1]).packed()
    with pytest.raises(nat.KsError):
        Trace(["a", "b"], [1, 1], [1, 1], [0, 9], [0, 0], [0, 1], [1, 1]).packed()
    with pytest.raises(nat.KsError):
        Trace(["a", "b"], [1, 1], [1, 1], [0, 0], [0, 0], [0, -1], [1, 1]).packed()
    empty = Trace([], [], [], [], [], [], [])
    assert empty.packed()[1].size >= 1


def test_cluster_pack(built_lib):
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, _native as nat
    c = ClusterSpec.load("gpu16")
    nconst, totals = c.packed()
    assert totals.tolist() == [1184000, 5898240, 64, 64000]
    rows = nconst.reshape(32, nat.KS_NC_FIELDS)   # node-major rows, padded to 32 nodes
    assert rows[:16, 0].tolist() == c.cpu_total.tolist()
    assert rows[:16, 2].tolist() == c.num_gpus.tolist()
    assert rows[:16, 4:12].tolist() == c.gpu_milli_total.tolist() and rows[:16, 12:20].tolist() == c.gpu_mem_total.tolist()
    assert not rows[16:].any()
    # slot 3 = node type: the index of the first node with an identical constant row (10 x P100 nodes, 2 x V100M32 ...)
    ty = rows[:16, 3].tolist()
    assert all(ty[i] <= i and rows[ty[i]].tolist()[:3] + rows[ty[i]].tolist()[4:] == rows[i].tolist()[:3] + rows[i].tolist()[4:]
               for i in range(16)) and len(set(ty)) == 5 and ty[0] == 0
    big = ClusterSpec.load("all1523").head(1024)
    assert big.n_pad == 1024 and big.packed()[1][2] == int(big.num_gpus.sum())


def test_plan_arithmetic(built_lib):
    from funsearch_kubernetes_simulator_b200 import _native as nat
    p = nat.KsPlan()
    p.n_episodes, p.n_nodes, p.max_pods, p.want_placements = 1024, 16, 8152, 0
    nat.check(nat.lib().ks_plan(C.byref(p)))
    assert p.warps_per_team == 1 and 1 <= p.teams_per_block <= 8 and (p.blocks * p.teams_per_block >= 1024 or p.blocks == 148)
    assert p.smem_per_block <= 232448 and p.heap_smem_entries == 0
    assert p.scratch_bytes >= nat.KS_SCRATCH_HEADER + p.blocks * p.teams_per_block * p.scratch_per_team
    q = nat.KsPlan()
    q.n_episodes, q.n_nodes, q.max_pods, q.want_placements = 3, 1024, 100, 1
    nat.check(nat.lib().ks_plan(C.byref(q)))
    assert q.blocks == 3 and q.teams_per_block == 1 and q.warps_per_team == 16 and q.smem_per_team >= 1024 * 48
    bad = nat.KsPlan()
    assert nat.lib().ks_plan(C.byref(bad)) != 0


def test_table_order_is_a_stable_policy_major_permutation():
    """evaluator.table_order: the permutation the evaluator applies to an episode table before upload."""
    from funsearch_kubernetes_simulator_b200.evaluator import table_order
    assert table_order(np.array([0, 0, 1, 2, 2])) is None and table_order(np.array([3])) is None
    pol = np.array([2, 0, 1, 0, 2, 1, 0])
    order = table_order(pol)
    assert order.tolist() == [1, 3, 6, 2, 5, 0, 4]          # by policy, ties in the caller's order
    inverse = np.argsort(order)
    assert (pol[order][inverse] == pol).all()               # results[inverse] is back in the caller's order


def test_plan_team_shapes(built_lib):
    """Narrower teams may fill the 512-thread kernel on request; two CTAs per SM only with one warp per episode."""
    from funsearch_kubernetes_simulator_b200 import _native as nat
    p = nat.KsPlan()
    p.n_episodes, p.n_nodes, p.max_pods, p.warps_per_team, p.teams_per_block = 296, 1024, 8152, 8, 2
    nat.check(nat.lib().ks_plan(C.byref(p)))
    assert p.teams_per_block == 2 and p.smem_per_block == 2 * p.smem_per_team <= 232448
    p.teams_per_block = 3
    assert nat.lib().ks_plan(C.byref(p)) != 0               # 24 warps do not fit a CTA
    q = nat.KsPlan()
    q.n_episodes, q.n_nodes, q.max_pods, q.warps_per_team, q.ctas_per_sm = 8192, 16, 8152, 2, 2
    assert nat.lib().ks_plan(C.byref(q)) != 0
    q.warps_per_team = 1
    nat.check(nat.lib().ks_plan(C.byref(q)))
    assert q.ctas_per_sm == 2 and q.blocks == 296 and q.teams_per_block == 8


def test_trace_pack_ranks_request_shapes(built_lib):
    """Slot 7 of a pod record = 1 + rank of the pod's (cpu, mem, gpus, gpu milli) among the trace's distinct request shapes,
    commonest first, ties by first appearance (the team kernels key their per-node score rows by it)."""
    from collections import Counter
    from funsearch_kubernetes_simulator_b200 import Trace, synthetic_trace
    for tr in (Trace.load("default"), Trace.load("gpushare60").head(3000), synthetic_trace(5000, seed=3, load_factor=200.0)):
        rank, podrec, _heap0, _place0 = tr.packed()
        shapes = list(zip(tr.cpu_milli.tolist(), tr.memory_mib.tolist(), tr.num_gpu.tolist(), tr.gpu_milli.tolist()))
        count = Counter(shapes)
        first = {}
        for i, s in enumerate(shapes):
            first.setdefault(s, i)
        order = sorted(count, key=lambda s: (-count[s], first[s]))
        want = {s: k + 1 for k, s in enumerate(order)}
        got = [int(podrec[int(rank[i]) * 8 + 7]) for i in range(tr.n_pods)]
        assert got == [want[s] for s in shapes], tr.name
        assert max(got) == len(count)


def test_team_plans_carry_score_rows_and_flags_mark_time_readers(built_lib):
    """Scratch of a team of warps holds 256 score rows (8 B x n_pad each) + their stamps on top of heap, placement records,
    waiting list and failure stamps; one-warp plans hold none.  The translator flags functions that read the pod's time fields."""
    from funsearch_kubernetes_simulator_b200 import _native as nat
    from funsearch_kubernetes_simulator_b200.policy_compiler import compile_policy
    from funsearch_kubernetes_simulator_b200.variants import PREFIX, SUFFIX
    solo, team = nat.KsPlan(), nat.KsPlan()
    solo.n_episodes, solo.n_nodes, solo.max_pods = 1024, 16, 8152
    team.n_episodes, team.n_nodes, team.max_pods = 64, 1024, 8152
    nat.check(nat.lib().ks_plan(C.byref(solo)))
    nat.check(nat.lib().ks_plan(C.byref(team)))
    base = 8152 * 8 * 2 + 8152 * 12          # heap + placement records + (waiting list, failure stamps)
    assert solo.warps_per_team == 1 and base <= solo.scratch_per_team < base + 1024
    rows = 256 * (4 + 8 * 1024)
    assert team.warps_per_team == 16 and base + rows <= team.scratch_per_team < base + rows + 1024
    flags = lambda body: compile_policy(PREFIX + body + SUFFIX).flags  # noqa: E731
    assert flags("    score = node.cpu_milli_left // 10\n") == nat.KS_PF_PREFIX_HOISTED
    assert flags("    score = pod.creation_time % 7 + 1\n") == nat.KS_PF_PREFIX_HOISTED | nat.KS_PF_READS_TIME
    assert flags("    score = pod.duration_time % 7 + 1\n") == nat.KS_PF_PREFIX_HOISTED | nat.KS_PF_READS_DURATION
    # a hand-written function that mentions the arguments gets the flags from the library itself
    src = 'extern "C" __device__ long long ks_hand_t(KS_POLICY_PARAMS) { return 1 + (p_ctime & 1) + (p_dur & 1); }\n'
    h = C.c_longlong(-1)
    nat.check(nat.lib().ks_policy_compile(src.encode(), b"ks_hand_t", 0, C.byref(h)))
    assert h.value >= 0
