"""GPU parity: the CUDA replay (through the C ABI) against the committed golden fixtures recorded
from the reference, and against the oracle on seeded inputs."""
import pytest

from helpers import assert_matches_golden, load_golden, specs_from_edge

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(built_lib):
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    gold, pol = load_golden()
    return dict(gold=gold, pol=pol, ClusterSpec=ClusterSpec, Trace=Trace, BatchEvaluator=BatchEvaluator)


def _run_group(env, cluster, trace_names_to_trace, tags):
    """tags: list of (tag, trace_name, policy_name); one launch for the whole group."""
    names = list(trace_names_to_trace)
    ev = env["BatchEvaluator"](cluster, [trace_names_to_trace[n] for n in names])
    pnames = sorted({p for _, _, p in tags})
    compiled = ev.compile([env["pol"][p] for p in pnames])
    assert all(c is not None for c in compiled)
    outs = ev.run_episodes(compiled, [pnames.index(p) for _, _, p in tags], [names.index(t) for _, t, _ in tags],
                           want_placements=True)
    for (tag, tname, _), out in zip(tags, outs):
        assert_matches_golden(tag, env["gold"]["runs"][tag], out, trace_names_to_trace[tname], cluster)
    return ev


def test_default_trace_shipped_policies(env):
    """BASELINE config 1: First-Fit, Best-Fit and the FunSearch champions on the bundled trace."""
    cluster = env["ClusterSpec"].load("gpu16")
    tr = env["Trace"].load("default")
    tags = [(t, "default", t.split("|")[2]) for t in env["gold"]["runs"] if t.startswith("default|gpu16|")]
    assert len(tags) >= 14
    _run_group(env, cluster, {"default": tr}, tags)


def test_all_bundled_traces_all_policies_one_launch(env):
    cluster = env["ClusterSpec"].load("gpu16")
    tags, traces = [], {}
    for t in env["gold"]["runs"]:
        parts = t.split("|")
        if len(parts) == 3 and parts[1] == "gpu16" and not t.startswith("edge:"):
            if parts[0] not in traces:
                traces[parts[0]] = env["Trace"].load(parts[0])
            tags.append((t, parts[0], parts[2]))
    assert len(traces) == 19
    _run_group(env, cluster, traces, tags)


def test_1024_node_cluster(env):
    cluster = env["ClusterSpec"].load("all1523").head(1024)
    tr = env["Trace"].load("default")
    tags = [(t, "default", t.split("|")[2]) for t in env["gold"]["runs"] if t.startswith("default|all1024|")]
    assert tags
    _run_group(env, cluster, {"default": tr}, tags)


def test_64_node_cluster_prefix(env):
    cluster = env["ClusterSpec"].load("gpu1213").head(64)
    tr = env["Trace"].load("cpu300").head(4000)
    tags = [(t, "x", t.split("|")[2]) for t in env["gold"]["runs"] if t.startswith("cpu300[:4000]|gpu64|")]
    assert tags
    _run_group(env, cluster, {"x": tr}, tags)


def test_edge_cases(env):
    for name, edge in env["gold"]["edge_cases"].items():
        cluster, trace = specs_from_edge(edge)
        tag = f"edge:{name}|{edge['policy']}"
        _run_group(env, cluster, {"e": trace}, [(tag, "e", edge["policy"])])


def test_launch_geometry_is_not_observable(env):
    """Episodes per CTA, and whether the placement records are an output or live in the team's scratch, must not
    change any result."""
    cluster = env["ClusterSpec"].load("gpu16")
    tr = env["Trace"].load("default")
    ev = env["BatchEvaluator"](cluster, [tr])
    pn = ["ts_funsearch_4901", "ts_first_fit"]
    compiled = ev.compile([env["pol"][p] for p in pn])
    for teams in (1, 2, 8):
        outs = ev.run_episodes(compiled, [0, 1], [0, 0], want_placements=True, teams_per_block=teams)
        assert ev.last_plan.teams_per_block == teams and ev.last_plan.heap_smem_entries == 0
        for p, out in zip(pn, outs):
            assert_matches_golden(f"teams={teams} {p}", env["gold"]["runs"][f"default|gpu16|{p}"], out, tr, cluster)
    outs = ev.run_episodes(compiled, [0, 1], [0, 0], want_placements=False)
    for p, out in zip(pn, outs):
        assert_matches_golden(f"no placements {p}", env["gold"]["runs"][f"default|gpu16|{p}"], out, tr, cluster)


def test_population_sample_against_oracle(env):
    """Seeded inputs the goldens do not cover: 24 generated policy variants on a 2,500-pod prefix, compared
    with the oracle run in this process (placement log, counters, metrics)."""
    from helpers import oracle_specs, outcome_digest
    from oracle import pysim
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    cluster = env["ClusterSpec"].load("gpu16")
    tr = env["Trace"].load("gpushare60").head(2500)
    pop = policy_population(24, seed=11)
    ev = env["BatchEvaluator"](cluster, [tr])
    compiled = ev.compile(pop)
    outs = ev.run_episodes(compiled, list(range(24)), [0] * 24, want_placements=True)
    cs, ts = oracle_specs(cluster, tr)
    for k, (src, out) in enumerate(zip(pop, outs)):
        ref = pysim.run_episode(cs, ts, pysim.scorer_from_source(src))
        assert (out.status, out.events, out.attempts, out.placed) == (ref.status, ref.events, ref.attempts, ref.placed), k
        assert out.max_nodes == ref.max_nodes and out.max_waiting == ref.max_waiting, k
        assert outcome_digest(tr, cluster, out) == pysim.placement_digest(ts, cs, ref), k
        assert abs(out.policy_score - ref.score) <= 1e-9 * max(abs(ref.score), 1e-300), k


def test_bench_population_on_the_full_trace_against_oracle(env):
    """The workload bench.py measures (configs[1]): the first 64 policies of the benchmarked generation on the FULL default
    trace, every counter, the placement log and all metrics against the oracle."""
    from helpers import oracle_specs, outcome_digest
    from oracle import pysim
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    cluster = env["ClusterSpec"].load("gpu16")
    tr = env["Trace"].load("default")
    pop = policy_population(1024, seed=0)[:64]
    ev = env["BatchEvaluator"](cluster, [tr])
    compiled = ev.compile(pop)
    assert all(c is not None for c in compiled)
    outs = ev.run_episodes(compiled, list(range(64)), [0] * 64, want_placements=True)
    cs, ts = oracle_specs(cluster, tr)
    declined = 0
    for k, (src, out) in enumerate(zip(pop, outs)):
        ref = pysim.run_episode(cs, ts, pysim.scorer_from_source(src))
        if out.status == 4:          # left the translated subset (e.g. a complex intermediate): reported, never a number
            declined += 1
            assert ref.status != 0 or out.policy_score == 0.0, k
            continue
        assert (out.status, out.events, out.attempts, out.placed) == (ref.status, ref.events, ref.attempts, ref.placed), k
        if ref.status != 0:
            assert out.policy_score == 0.0, k
            continue
        assert (out.max_nodes, out.max_waiting, out.last_event_time) == (ref.max_nodes, ref.max_waiting, ref.last_event_time), k
        assert outcome_digest(tr, cluster, out) == pysim.placement_digest(ts, cs, ref), k
        r = out.results
        assert (r.num_snapshots, r.num_fragmentation_events) == (ref.num_snapshots, ref.num_fragmentation_events), k
        for got, want in ((r.avg_cpu_utilization, ref.avg_cpu), (r.avg_memory_utilization, ref.avg_mem),
                          (r.avg_gpu_count_utilization, ref.avg_gpu_count), (r.avg_gpu_memory_utilization, ref.avg_gpu_milli),
                          (r.gpu_fragmentation_score, ref.fragmentation), (out.policy_score, ref.score)):
            assert abs(got - want) <= 1e-9 * max(abs(want), 1e-300), (k, got, want)
    assert declined <= 3


def test_bootstrap_seeds_on_full_traces_against_oracle(env):
    """configs[3]'s "seeds": bootstrap resamples 1-3 of two FULL bundled traces (one pre-sorted, one not) under three
    policies, against the oracle."""
    from helpers import oracle_specs, outcome_digest
    from oracle import pysim
    from funsearch_kubernetes_simulator_b200.workload import bootstrap_trace
    cluster = env["ClusterSpec"].load("gpu16")
    traces = [bootstrap_trace(env["Trace"].load(nm), s) for nm in ("default", "cpu200") for s in (1, 2, 3)]
    pn = ["ts_best_fit", "ts_funsearch_4816", "fs_utilization_based"]
    ev = env["BatchEvaluator"](cluster, traces)
    compiled = ev.compile([env["pol"][p] for p in pn])
    ep_pol = [p for p in range(3) for _t in range(6)]
    ep_tr = [t for _p in range(3) for t in range(6)]
    outs = ev.run_episodes(compiled, ep_pol, ep_tr, want_placements=True)
    for e, out in enumerate(outs):
        tr = traces[ep_tr[e]]
        cs, ts = oracle_specs(cluster, tr)
        ref = pysim.run_episode(cs, ts, pysim.scorer_from_source(env["pol"][pn[ep_pol[e]]]))
        assert (out.status, out.events, out.attempts, out.placed) == (ref.status, ref.events, ref.attempts, ref.placed), e
        assert (out.max_nodes, out.max_waiting, out.last_event_time) == (ref.max_nodes, ref.max_waiting, ref.last_event_time), e
        assert outcome_digest(tr, cluster, out) == pysim.placement_digest(ts, cs, ref), e
        assert abs(out.policy_score - ref.score) <= 1e-9 * max(abs(ref.score), 1e-300), e


def test_policy_major_table_and_two_ctas_per_sm(env):
    """A trace-major episode table is replayed policy-major (same-policy episodes side by side on an SM) and, being
    several waves long with many episodes per policy, by the two-CTAs-per-SM kernel; results come back in the caller's
    order, identical to the one-CTA kernel's and to the oracle's."""
    import numpy as np
    from helpers import oracle_specs
    from oracle import pysim
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    from funsearch_kubernetes_simulator_b200.workload import bootstrap_trace
    cluster = env["ClusterSpec"].load("gpu16")
    base = env["Trace"].load("default").head(250)
    traces = [bootstrap_trace(base, s) for s in range(128)]
    pop = policy_population(37, seed=5)
    ev = env["BatchEvaluator"](cluster, traces)
    compiled = ev.compile(pop)
    ep_pol = [p for t in range(128) for p in range(37)]
    ep_tr = [t for t in range(128) for p in range(37)]
    rb = ev.prepare_resident(compiled, ep_pol, ep_tr)
    assert rb.plan.ctas_per_sm == 2 and rb.order is not None and rb.plan.blocks > ev.sm_count
    rb.launch()
    dense = rb.results()
    assert (dense["status"] == 0).all()
    ev.ctas_per_sm = 1
    rb1 = ev.prepare_resident(compiled, ep_pol, ep_tr)
    assert rb1.plan.ctas_per_sm == 1
    rb1.launch()
    one = rb1.results()
    for f in dense.dtype.names:
        if f not in ("device_cycles", "reserved"):
            assert np.array_equal(dense[f], one[f]), f
    scores = rb.scores_tensor().cpu().numpy()
    assert np.array_equal(scores, dense["policy_score"])
    hb = ev.prepare_host_batch(compiled, ep_pol, ep_tr)
    assert np.array_equal(hb.run()["events"], dense["events"])
    for e in (0, 36, 37 * 5 + 3, 37 * 127 + 36):   # the caller's order: episode e = (policy e % 37, trace e // 37)
        cs, ts = oracle_specs(cluster, traces[ep_tr[e]])
        ref = pysim.run_episode(cs, ts, pysim.scorer_from_source(pop[ep_pol[e]]))
        assert (dense["events"][e], dense["attempts"][e], dense["placed"][e]) == (ref.events, ref.attempts, ref.placed), e
        assert abs(dense["policy_score"][e] - ref.score) <= 1e-9 * max(abs(ref.score), 1e-300), e


def test_batch_spread_over_several_modules_equals_one_module(env):
    """A generation whose structures live in several linked modules (what the compile workers produce: one module per
    worker) is replayed by one launch per module, in flight together; every result equals the single-module replay and
    the oracle, through all three call forms."""
    import numpy as np
    from helpers import oracle_specs
    from oracle import pysim
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    cluster = env["ClusterSpec"].load("gpu16")
    traces = [env["Trace"].load("default").head(900), env["Trace"].load("cpu200").head(700)]
    pop = policy_population(40, seed=17)
    ev = env["BatchEvaluator"](cluster, traces)
    ev.module_structures = 5
    compiled = ev.compile(pop, processes=0)
    ep_pol = [p for p in range(40) for _ in range(2)]
    ep_tr = [t for _ in range(40) for t in range(2)]
    rb = ev.prepare_resident(compiled, ep_pol, ep_tr)
    assert rb.n_modules >= 4 and len({p.teams_per_block for p in rb.plans}) == 1
    rb.launch()
    many = rb.results()
    ev1 = env["BatchEvaluator"](cluster, traces)
    rb1 = ev1.prepare_resident(ev1.compile(pop, processes=0), ep_pol, ep_tr)
    assert rb1.n_modules == 1
    rb1.launch()
    one = rb1.results()
    for f in many.dtype.names:
        if f not in ("device_cycles", "reserved"):
            assert np.array_equal(many[f], one[f]), f
    assert np.array_equal(rb.scores_tensor().cpu().numpy(), one["policy_score"])
    hb = ev.prepare_host_batch(compiled, ep_pol, ep_tr)
    assert len(hb.calls) == rb.n_modules
    host = hb.run()
    for f in ("status", "events", "attempts", "placed", "policy_score"):
        assert np.array_equal(host[f], one[f]), f
    outs = ev.run_episodes(compiled, ep_pol, ep_tr, want_placements=True)
    from helpers import outcome_digest
    for e in (0, 7, 38, 79):
        cs, ts = oracle_specs(cluster, traces[ep_tr[e]])
        ref = pysim.run_episode(cs, ts, pysim.scorer_from_source(pop[ep_pol[e]]))
        assert (outs[e].status, outs[e].events, outs[e].attempts, outs[e].placed) == (ref.status, ref.events, ref.attempts, ref.placed), e
        if ref.status == 0:
            assert outcome_digest(traces[ep_tr[e]], cluster, outs[e]) == pysim.placement_digest(ts, cs, ref), e
        assert abs(outs[e].policy_score - ref.score) <= 1e-9 * max(abs(ref.score), 1e-300), e


def test_generation_compiled_and_linked_by_the_worker_pool(env):
    """A fresh generation through evaluate_population with the worker pool: translation, NVRTC and linking happen in the
    workers (one module per worker), this process loads the modules; scores equal the in-process path's."""
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    cluster = env["ClusterSpec"].load("gpu16")
    tr = env["Trace"].load("default").head(600)
    pop = policy_population(96, seed=23)
    ev = env["BatchEvaluator"](cluster, [tr], processes=4)
    got = ev.evaluate_population(pop)
    st = ev.last_compile_stats
    assert st["pool_workers"] >= 2 and 2 <= st["modules_loaded"] <= 4 and st["new_structures"] >= 48, st
    ev1 = env["BatchEvaluator"](cluster, [tr], processes=0)
    from funsearch_kubernetes_simulator_b200 import policy_compiler
    want = ev1.evaluate_population(pop)
    assert got == want and sum(1 for x in got if x > 0) > 60


def test_host_buffer_entry_point_matches_resident_path(env):
    """ks_eval_batch_host (pinned host buffers) == ks_eval_batch (buffers resident in HBM)."""
    import numpy as np
    cluster = env["ClusterSpec"].load("gpu16")
    tr = [env["Trace"].load("default").head(3000), env["Trace"].load("cpu100").head(1777)]
    ev = env["BatchEvaluator"](cluster, tr)
    pn = ["ts_funsearch_4901", "ts_best_fit", "fs_gpu_aware"]
    compiled = ev.compile([env["pol"][p] for p in pn])
    ep_pol, ep_tr = [0, 1, 2, 0, 1, 2], [0, 0, 0, 1, 1, 1]
    rb = ev.prepare_resident(compiled, ep_pol, ep_tr)
    rb.launch()
    a = rb.results()
    hb = ev.prepare_host_batch(compiled, ep_pol, ep_tr)
    b = hb.run()
    assert hb.h2d_bytes > 0 and hb.d2h_bytes == 6 * 128
    for f in a.dtype.names:
        if f not in ("device_cycles", "reserved"):
            assert np.array_equal(a[f], b[f]), f
    outs = ev.run_episodes(compiled, ep_pol, ep_tr)
    assert [o.events for o in outs] == a["events"].tolist()


def test_evaluate_population_reference_shaped_api(env):
    """evaluate_population / evaluate_policy_standalone: scores as the reference's worker returns them,
    0 for policies that raise or do not translate (funsearch_integration.py:63-64)."""
    cluster = env["ClusterSpec"].load("gpu16")
    tr = env["Trace"].load("default")
    ev = env["BatchEvaluator"](cluster, [tr])
    codes = [env["pol"]["ts_first_fit"], env["pol"]["ts_best_fit"], "def priority_function(pod, node):\n    return 1 / 0\n",
             "def priority_function(pod, node):\n    import os\n    return 1\n", "this is not python"]
    scores = ev.evaluate_population(codes)
    want = [float.fromhex(env["gold"]["runs"][f"default|gpu16|{p}"]["score"]) for p in ("ts_first_fit", "ts_best_fit")]
    assert abs(scores[0] - want[0]) < 1e-12 and abs(scores[1] - want[1]) < 1e-12
    assert scores[2:] == [0.0, 0.0, 0.0]
    idx, code, s = ev.evaluate_policy_standalone((7, codes[1]))
    assert idx == 7 and code == codes[1] and abs(s - want[1]) < 1e-12


@pytest.mark.parametrize("width", [2, 4, 8, 16])
def test_multi_warp_teams_give_identical_results(env, width):
    """An episode replayed by a team of W warps (leader + helpers at a named barrier) must equal W = 1."""
    # 16 nodes: most helper lanes own no node at all
    cluster = env["ClusterSpec"].load("gpu16")
    tr = env["Trace"].load("default")
    ev = env["BatchEvaluator"](cluster, [tr])
    ev.warps_per_team = width
    pn = ["ts_funsearch_4901", "ts_first_fit", "fs_worst_fit", "x_ignores_capacity"]
    compiled = ev.compile([env["pol"][p] for p in pn])
    outs = ev.run_episodes(compiled, list(range(len(pn))), [0] * len(pn), want_placements=True)
    assert ev.last_plan.warps_per_team == width
    for p, out in zip(pn, outs):
        assert_matches_golden(f"W={width} {p}", env["gold"]["runs"][f"default|gpu16|{p}"], out, tr, cluster)
    # 1,024 nodes: every lane owns nodes
    big = env["ClusterSpec"].load("all1523").head(1024)
    ev = env["BatchEvaluator"](big, [tr])
    ev.warps_per_team = width
    pn = ["ts_best_fit", "ts_funsearch_4816"]
    compiled = ev.compile([env["pol"][p] for p in pn])
    outs = ev.run_episodes(compiled, [0, 1], [0, 0], want_placements=True)
    for p, out in zip(pn, outs):
        assert_matches_golden(f"W={width} 1024 nodes {p}", env["gold"]["runs"][f"default|all1024|{p}"], out, tr, big)
    # 64 nodes with retries and a busy waiting list
    c64 = env["ClusterSpec"].load("gpu1213").head(64)
    tr2 = env["Trace"].load("cpu300").head(4000)
    ev = env["BatchEvaluator"](c64, [tr2])
    ev.warps_per_team = width
    pn = ["ts_best_fit", "ts_funsearch_4901", "ts_funsearch_4800"]
    compiled = ev.compile([env["pol"][p] for p in pn])
    outs = ev.run_episodes(compiled, [0, 1, 2], [0, 0, 0], want_placements=True)
    for p, out in zip(pn, outs):
        assert_matches_golden(f"W={width} 64 nodes {p}", env["gold"]["runs"][f"cpu300[:4000]|gpu64|{p}"], out, tr2, c64)


def _compare_with_c_oracle(tag, cluster, trace, out, builtin, params=None):
    from oracle import coracle
    res, t, n, m = coracle.run(cluster, trace, builtin, params)
    assert (out.status, out.events, out.attempts, out.placed) == (res.status, res.events, res.attempts, res.placed), tag
    assert (out.max_nodes, out.max_waiting, out.last_event_time) == (res.max_nodes, res.max_waiting, res.last_event_time), tag
    ns = out.results.num_snapshots if out.results else 0
    assert ns == res.num_snapshots and (out.results.num_fragmentation_events if out.results else 0) == res.num_frag, tag
    import numpy as np
    assert np.array_equal(out.final_creation_time, t) and np.array_equal(out.node_index, n) and np.array_equal(out.gpu_mask, m), tag
    for got, want in ((out.policy_score, res.score), (out.results.avg_cpu_utilization, res.avg_cpu),
                      (out.results.avg_memory_utilization, res.avg_mem), (out.results.avg_gpu_count_utilization, res.avg_gpu_count),
                      (out.results.avg_gpu_memory_utilization, res.avg_gpu_milli),
                      (out.results.gpu_fragmentation_score, res.fragmentation)):
        assert abs(got - want) <= 1e-9 * max(abs(want), 1e-300), (tag, got, want)


BUILTIN = {"ts_first_fit": "first_fit", "ts_best_fit": "best_fit", "ts_funsearch_4901": "funsearch_4901",
           "ts_funsearch_4816": "funsearch_4816", "ts_funsearch_4800": "funsearch_4800", "fs_worst_fit": "worst_fit"}


def test_config3_shape_100k_pods_1024_nodes(env):
    """BASELINE config 3, scaled to what the C oracle replays in seconds: synthetic Alibaba-shaped trace on the
    first 1,024 rows of the all-node list, heavy enough that retries occur."""
    from funsearch_kubernetes_simulator_b200 import synthetic_trace
    cluster = env["ClusterSpec"].load("all1523").head(1024)
    trace = synthetic_trace(100000, seed=1, load_factor=1500.0)
    ev = env["BatchEvaluator"](cluster, [trace])
    names = list(BUILTIN)
    compiled = ev.compile([env["pol"][p] for p in names])
    outs = ev.run_episodes(compiled, list(range(len(names))), [0] * len(names), want_placements=True)
    assert ev.last_plan.warps_per_team == 16
    assert any(o.results.num_fragmentation_events > 1000 for o in outs)
    for p, out in zip(names, outs):
        _compare_with_c_oracle(p, cluster, trace, out, BUILTIN[p])


def test_config3_full_size_1m_pods_1024_nodes(env):
    """BASELINE config 3 at full size (1,000,000 pods x 1,024 nodes, load factor 200: the cluster just keeps
    up under a packing policy): three shipped policies bit-exact against the C oracle, plus size-independent
    invariants on generated policy variants (under an event budget: a spreading policy never catches up)."""
    import threading
    import numpy as np
    from funsearch_kubernetes_simulator_b200 import synthetic_trace
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    from oracle import coracle
    P, BUDGET = 1000000, 4000000
    cluster = env["ClusterSpec"].load("all1523").head(1024)
    trace = synthetic_trace(P, seed=3, load_factor=200.0)
    names = ["ts_first_fit", "ts_best_fit", "ts_funsearch_4816"]
    ref = {}

    def work(p):
        ref[p] = coracle.run(cluster, trace, BUILTIN[p], max_events=BUDGET)   # ctypes releases the GIL
    threads = [threading.Thread(target=work, args=(p,)) for p in names]
    for th in threads:
        th.start()
    variants = policy_population(3, seed=21)
    ev = env["BatchEvaluator"](cluster, [trace])
    compiled = ev.compile([env["pol"][p] for p in names] + variants)
    outs = ev.run_episodes(compiled, list(range(len(compiled))), [0] * len(compiled), want_placements=True, max_events=BUDGET)
    for th in threads:
        th.join()
    for p, out in zip(names, outs):
        res, t, n, m = ref[p]
        assert res.status == 0 and res.placed == P, p
        assert (out.status, out.events, out.attempts, out.placed) == (res.status, res.events, res.attempts, res.placed), p
        assert np.array_equal(out.final_creation_time, t) and np.array_equal(out.node_index, n) and np.array_equal(out.gpu_mask, m), p
        assert abs(out.policy_score - res.score) <= 1e-9 * max(abs(res.score), 1e-300), p
    for k, out in enumerate(outs):
        nf = out.results.num_fragmentation_events
        assert out.status in (0, 3) and out.attempts == out.placed + nf, k
        if out.status == 0:
            assert out.events == out.attempts + out.placed and out.placed == P and (out.node_index >= 0).all(), k
        else:
            assert out.events == BUDGET and out.policy_score == 0.0, k
        assert int((out.node_index >= 0).sum()) == out.placed, k
        # snapshots: thresholds 0.05, 0.1, ... (fp64 accumulation) not above events / P
        thr, cnt = 0.05, 0
        while out.events / P >= thr:
            cnt += 1
            thr += 0.05
        assert out.results.num_snapshots == cnt, k
        assert 0.0 <= out.policy_score <= 1.0


def test_object_protocol_facade(env):
    """GpuKubernetesSimulator: the reference's run_schedule() protocol on entity objects (tests/test_simulator.py
    scenario: known answer pod1->node2, pod2->node1 [0], pod3->node2, pod4->node1 [0,1])."""
    from types import SimpleNamespace as NS
    from funsearch_kubernetes_simulator_b200.integration import GpuKubernetesSimulator

    def gpu():
        return NS(memory_mib_left=8000, memory_mib_total=8000, gpu_milli_left=1000, gpu_milli_total=1000)
    node1 = NS(node_id="node1", cpu_milli_left=8000, cpu_milli_total=8000, memory_mib_left=16000, memory_mib_total=16000,
               gpu_left=2, gpus=[gpu(), gpu()])
    node2 = NS(node_id="node2", cpu_milli_left=4000, cpu_milli_total=4000, memory_mib_left=8000, memory_mib_total=8000,
               gpu_left=0, gpus=[])
    cluster = NS(nodes_dict={"node1": node1, "node2": node2})
    rows = [("pod1", 1000, 2000, 0, 0, 0, 10), ("pod2", 2000, 4000, 1, 500, 5, 15), ("pod3", 3000, 6000, 0, 0, 10, 8),
            ("pod4", 1500, 3000, 2, 400, 15, 12)]
    pods = [NS(pod_id=r[0], cpu_milli=r[1], memory_mib=r[2], num_gpu=r[3], gpu_milli=r[4], gpu_spec="", creation_time=r[5],
               duration_time=r[6], assigned_node="", assigned_gpus=[]) for r in rows]
    sim = GpuKubernetesSimulator(cluster, pods, env["pol"]["sim_int_best_fit"], validate_invariants=True)
    sim.run_schedule()
    assert [p.assigned_node for p in pods] == ["node2", "node1", "node2", "node1"]
    assert [p.assigned_gpus for p in pods] == [[], [0], [], [0, 1]]
    g = env["gold"]["runs"]["edge:test_simulator|sim_int_best_fit"]
    assert abs(sim.get_policy_score() - float.fromhex(g["score"])) < 1e-12 and sim.max_nodes == g["max_nodes"]
    assert sim.get_evaluation_results().num_snapshots == g["num_snapshots"]
    # the rest of the object protocol: nothing is left waiting, every resource is back at its total (main.py:83-96)
    assert sim.waiting_pods == [] and (node1.cpu_milli_left, node1.gpu_left, node1.gpus[1].gpu_milli_left) == (8000, 2, 1000)
    # an oversized pod with no deletion pending is dropped and stays in waiting_pods (SURVEY App. A.9 case 1)
    big = NS(pod_id="pod9", cpu_milli=99000, memory_mib=1, num_gpu=0, gpu_milli=0, gpu_spec="", creation_time=1,
             duration_time=5, assigned_node="", assigned_gpus=[])
    pods2 = [NS(**vars(p)) for p in pods[:1]] + [big]
    for p in pods2:
        p.assigned_node, p.assigned_gpus = "", []
    pods2[0].creation_time = 50
    sim2 = GpuKubernetesSimulator(cluster, pods2, env["pol"]["sim_int_best_fit"])
    sim2.run_schedule()
    assert sim2.waiting_pods == [big] and big.assigned_node == "" and pods2[0].assigned_node == "node2"
    assert sim2.get_policy_score() == 0.0
    node1.cpu_milli_left -= 1
    with pytest.raises(ValueError, match="not pristine"):
        GpuKubernetesSimulator(cluster, pods, env["pol"]["sim_int_best_fit"])
    node1.cpu_milli_left += 1


def test_device_side_invariant_checks_in_the_debug_build(env):
    """validate_invariants=True on the device: the debug build of the library (build.py --invariants) runs the reference's
    _validate_cluster_invariants (main.py:201-272) on the kernel's live state after every mutation.  A policy that ignores
    capacity drives a node negative and the episode stops there with the reference's message; well-behaved policies
    replay exactly as in the production build.  Run in a subprocess: the library variant is chosen at import time."""
    import os
    import subprocess
    import sys
    from funsearch_kubernetes_simulator_b200 import build
    lib = build.build(invariants=True)
    script = r'''
import json, sys
from types import SimpleNamespace as NS
sys.path.insert(0, "tests")
from helpers import load_golden
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, _native as nat
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.integration import GpuKubernetesSimulator
assert nat.lib().ks_build_flags() & 2
gold, pol = load_golden()
c, t = ClusterSpec.load("gpu16"), Trace.load("default").head(3000)
def objects():
    nodes = {}
    for i, nid in enumerate(c.node_ids):
        gpus = [NS(memory_mib_left=int(c.gpu_mem_total[i, g]), memory_mib_total=int(c.gpu_mem_total[i, g]),
                   gpu_milli_left=int(c.gpu_milli_total[i, g]), gpu_milli_total=int(c.gpu_milli_total[i, g])) for g in range(int(c.num_gpus[i]))]
        nodes[nid] = NS(node_id=nid, cpu_milli_left=int(c.cpu_total[i]), cpu_milli_total=int(c.cpu_total[i]),
                        memory_mib_left=int(c.mem_total[i]), memory_mib_total=int(c.mem_total[i]), gpu_left=len(gpus), gpus=gpus)
    pods = [NS(pod_id=t.pod_ids[i], cpu_milli=int(t.cpu_milli[i]), memory_mib=int(t.memory_mib[i]), num_gpu=int(t.num_gpu[i]),
               gpu_milli=int(t.gpu_milli[i]), gpu_spec="", creation_time=int(t.creation_time[i]), duration_time=int(t.duration_time[i]),
               assigned_node="", assigned_gpus=[]) for i in range(t.n_pods)]
    return NS(nodes_dict=nodes), pods
IGNORES_CPU_AND_MEMORY = """
def priority_function(pod, node):
    if pod.num_gpu > node.gpu_left:
        return 0
    if pod.num_gpu > 0:
        n = 0
        for g in node.gpus:
            if g.gpu_milli_left >= pod.gpu_milli:
                n += 1
        if n < pod.num_gpu:
            return 0
    return 1000 - len(node.gpus)
"""
out = {}
cl, pods = objects()
try:
    GpuKubernetesSimulator(cl, pods, IGNORES_CPU_AND_MEMORY, validate_invariants=True).run_schedule()
    out["bad"] = "no error"
except ValueError as e:
    out["bad"] = str(e)
cl, pods = objects()
sim = GpuKubernetesSimulator(cl, pods, pol["ts_funsearch_4816"], validate_invariants=True)
sim.run_schedule()
out["good"] = [sim.outcome.events, sim.outcome.placed, sim.get_policy_score().hex()]
print("RESULT " + json.dumps(out))
'''
    root = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
    r = subprocess.run([sys.executable, "-c", script], cwd=root, env=dict(os.environ, KS_B200_LIB=lib, PYTHONPATH=root),
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    import json
    out = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")][-1][7:])
    assert "has negative CPU remaining" in out["bad"] or "has negative memory remaining" in out["bad"], out["bad"]
    assert out["bad"].startswith("Node openb-node-")
    # the same episode in the production build of this process
    cluster, tr = env["ClusterSpec"].load("gpu16"), env["Trace"].load("default").head(3000)
    ev = env["BatchEvaluator"](cluster, [tr])
    o = ev.run_episodes(ev.compile([env["pol"]["ts_funsearch_4816"]]), [0], [0])[0]
    assert out["good"] == [o.events, o.placed, o.policy_score.hex()]


def test_device_pow_is_bit_identical_to_host_libm(env):
    """`**` in a policy is libm pow() in CPython; the device must return the same bits (glibc-exact fast path).
    The reference value is computed by this process's own libm, i.e. the GPU box's host: same image, same glibc."""
    import ctypes as C
    import math
    import numpy as np
    import torch
    from funsearch_kubernetes_simulator_b200 import _native as nat
    ev = env["BatchEvaluator"](env["ClusterSpec"].load("gpu16"), [env["Trace"].load("default").head(10)])
    compiled = ev.compile([env["pol"]["ts_first_fit"]])
    prog = ev._program((compiled[0].handle,))
    rng = np.random.default_rng(12)
    n = 400000
    x = np.concatenate([rng.random(n // 4), np.exp((rng.random(n // 4) - 0.3) * 40.0),
                        rng.integers(1, 2000000, n // 4).astype(np.float64), rng.random(n // 4) * 1000 + 1e-9])
    ys = np.array([0.5, 0.6, 2.0, 2.5, 3.0, 1.5, 0.25, -1.0, -0.5, 1e-3, 7.0, 1 / 3])
    y = np.concatenate([ys[rng.integers(0, 12, n // 4)], (rng.random(n // 4) - 0.5) * 20, ys[rng.integers(0, 12, n // 4)],
                        rng.random(n // 4) * 4])
    # outside glibc's fast path: subnormal x, tiny and huge |y|, results through overflow and underflow
    m = 50000
    x2 = np.concatenate([rng.integers(1, 2 ** 52, m).astype(np.uint64).view(np.float64), np.exp((rng.random(m) - 0.5) * 8),
                         np.exp((rng.random(m) - 0.5) * 8), np.exp((rng.random(m) - 0.5) * 1400)])
    y2 = np.concatenate([(rng.random(m) - 0.5) * 3, np.ldexp(rng.random(m) - 0.5, rng.integers(-110, -50, m)),
                         np.ldexp(rng.random(m) - 0.5, rng.integers(62, 92, m)), (rng.random(m) - 0.5) * 8])
    keep = y2 != 0.0
    x, y = np.concatenate([x, x2[keep]]), np.concatenate([y, y2[keep]])
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    out = torch.empty_like(dx)
    nat.check(nat.lib().ks_debug_pow(prog, C.c_void_p(dx.data_ptr()), C.c_void_p(dy.data_ptr()), C.c_void_p(out.data_ptr()),
                                     len(x), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    got = out.cpu().numpy()

    def host(a, b):
        try:
            return math.pow(a, b)
        except OverflowError:
            return math.inf
    want = np.array([host(a, b) for a, b in zip(x.tolist(), y.tolist())])
    # the one band that is not ported is declined (a tagged NaN -> status 'unsupported'), never approximated
    declined = (got.view(np.uint64) >> np.uint64(32)) == np.uint64(0x7ff8dead)
    band = np.abs(y * np.log(x))
    assert ((band[declined] > 500) & (band[declined] < 1030)).all() and 0 < declined.sum() <= m and not declined[:n].any()
    bad = np.nonzero((got.view(np.uint64) != want.view(np.uint64)) & ~declined)[0]
    assert bad.size == 0, f"{bad.size} of {len(x)} differ, e.g. x={x[bad[0]]!r} y={y[bad[0]]!r} got={got[bad[0]]!r} want={want[bad[0]]!r}"


def test_device_division_is_ieee_division(env):
    """ks_rt_ddiv (leaf fast path + out-of-line rare operands) must be IEEE-754 division bit for bit: operands of every
    scale, subnormals, zeros, infinities, NaNs, quotients that overflow or underflow."""
    import ctypes as C
    import numpy as np
    import torch
    from funsearch_kubernetes_simulator_b200 import _native as nat
    ev = env["BatchEvaluator"](env["ClusterSpec"].load("gpu16"), [env["Trace"].load("default").head(10)])
    compiled = ev.compile([env["pol"]["ts_first_fit"]])
    prog = ev._program((compiled[0].handle,))
    rng = np.random.default_rng(21)
    n = 1 << 20
    q = n // 8
    bits = lambda k: rng.integers(0, 2 ** 64, k, dtype=np.uint64).view(np.float64)           # any bit pattern
    ints = lambda k: rng.integers(-2 ** 40, 2 ** 40, k).astype(np.float64)
    wide = lambda k: np.ldexp(rng.random(k) + 0.5, rng.integers(-1070, 1020, k)) * rng.choice([-1.0, 1.0], k)
    edge = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308,
                     2.0 ** -255, 2.0 ** 257, 2.0 ** -256, 2.0 ** 256, 3.0, 1 / 3])
    x = np.concatenate([bits(q), ints(q), wide(q), rng.random(q), ints(q), wide(q), rng.random(q) * 1e6, np.repeat(edge, q // 16)])
    y = np.concatenate([bits(q), ints(q), wide(q), rng.random(q), rng.random(q), ints(q), ints(q), np.tile(edge, q // 16)])
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    out = torch.empty_like(dx)
    nat.check(nat.lib().ks_debug_math(prog, 3, C.c_void_p(dx.data_ptr()), C.c_void_p(dy.data_ptr()), C.c_void_p(out.data_ptr()),
                                      len(x), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    got = out.cpu().numpy()
    with np.errstate(all="ignore"):
        want = x / y
    same = (got.view(np.uint64) == want.view(np.uint64)) | (np.isnan(got) & np.isnan(want))
    bad = np.nonzero(~same)[0]
    assert bad.size == 0, f"{bad.size} of {len(x)} differ, e.g. {x[bad[0]]!r} / {y[bad[0]]!r}: got {got[bad[0]]!r} want {want[bad[0]]!r}"


def test_device_exp_log_bit_identical_to_host_libm(env):
    """math.exp / math.log in a policy are glibc's exp and log; the device versions (csrc/ks_glibc_explog.h) must return
    the same bits as this host's libm over the whole domain, overflow, subnormal results and subnormal arguments included."""
    import ctypes as C
    import math
    import numpy as np
    import torch
    from funsearch_kubernetes_simulator_b200 import _native as nat
    ev = env["BatchEvaluator"](env["ClusterSpec"].load("gpu16"), [env["Trace"].load("default").head(10)])
    compiled = ev.compile([env["pol"]["ts_first_fit"]])
    prog = ev._program((compiled[0].handle,))
    rng = np.random.default_rng(13)
    n = 200000
    q = n // 5

    def run(op, x):
        dx = torch.from_numpy(x).cuda()
        out = torch.empty_like(dx)
        nat.check(nat.lib().ks_debug_math(prog, op, C.c_void_p(dx.data_ptr()), C.c_void_p(dx.data_ptr()), C.c_void_p(out.data_ptr()),
                                          len(x), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return out.cpu().numpy()

    def host(fn, x):
        out = np.empty_like(x)
        for i, v in enumerate(x.tolist()):
            try:
                out[i] = fn(v)
            except OverflowError:
                out[i] = math.inf
        return out

    xe = np.concatenate([(rng.random(q) - 0.5) * 4, (rng.random(q) - 0.5) * 1500, -700 - rng.random(q) * 50,
                         (rng.random(q) - 0.5) * 1e-10, 700 + rng.random(q) * 12, np.array([0.0, -0.0, math.inf, -math.inf, 709.782712893384, -745.2])])
    got, want = run(1, xe), host(math.exp, xe)
    bad = np.nonzero(got.view(np.uint64) != want.view(np.uint64))[0]
    assert bad.size == 0, f"exp: {bad.size} differ, e.g. x={xe[bad[0]]!r} got={got[bad[0]]!r} want={want[bad[0]]!r}"
    xl = np.concatenate([0.9 + rng.random(q) * 0.2, np.exp((rng.random(q) - 0.5) * 1400), rng.integers(1, 2000000, q).astype(np.float64),
                         rng.integers(1, 2 ** 52, q).astype(np.uint64).view(np.float64),
                         rng.random(q) * 1000 + 1e-300, np.array([1.0, math.inf, 5e-324, 2.2250738585072014e-308])])
    got, want = run(2, xl), host(math.log, xl)
    bad = np.nonzero(got.view(np.uint64) != want.view(np.uint64))[0]
    assert bad.size == 0, f"log: {bad.size} differ, e.g. x={xl[bad[0]]!r} got={got[bad[0]]!r} want={want[bad[0]]!r}"


# ---- retries: the kernel rescans only the nodes that changed since the pod last failed, unless the policy sees the time ----
def _retry_policies():
    from funsearch_kubernetes_simulator_b200.variants import PREFIX, SUFFIX
    return {
        # pure in (pod fields, node rows): the shortcut applies
        "prefixed_pure": PREFIX + "    score = 1000 - node.cpu_milli_left // 100 + node.gpu_left * 3\n" + SUFFIX,
        # reads pod.creation_time, which every retry changes: a pod refused for its time alone gets in on a later retry
        # although no node changed in between, so the shortcut must be off
        "prefixed_reads_time": PREFIX + "    if pod.creation_time % 5 < 3:\n        return 0\n"
                                        "    score = 500 + node.memory_mib_left // 1000\n" + SUFFIX,
        # reads pod.duration_time: pods of one request shape score differently, so no scores are kept per shape (teams)
        "prefixed_reads_duration": PREFIX + "    score = 100 + (pod.duration_time % 97) * (node.gpu_left + 1) + node.cpu_milli_left // 500\n" + SUFFIX,
        # no template prefix: the function is its own feasibility test and is called for every (changed) node
        "raw_own_feasibility": "def priority_function(pod, node):\n"
                               "    if pod.cpu_milli > node.cpu_milli_left or pod.memory_mib > node.memory_mib_left:\n        return 0\n"
                               "    if pod.num_gpu > node.gpu_left:\n        return 0\n"
                               "    free = 0\n"
                               "    for g in node.gpus:\n"
                               "        if g.gpu_milli_left >= pod.gpu_milli:\n            free += 1\n"
                               "    if free < pod.num_gpu:\n        return 0\n"
                               "    return 1 + node.cpu_milli_left // 1000\n",
        # raw and time-dependent
        "raw_reads_time": "def priority_function(pod, node):\n"
                          "    if pod.cpu_milli > node.cpu_milli_left or pod.memory_mib > node.memory_mib_left or pod.num_gpu > node.gpu_left:\n"
                          "        return 0\n"
                          "    free = 0\n"
                          "    for g in node.gpus:\n"
                          "        if g.gpu_milli_left >= pod.gpu_milli:\n            free += 1\n"
                          "    if free < pod.num_gpu or (pod.creation_time + pod.duration_time) % 3 == 0:\n        return 0\n"
                          "    return 7\n",
    }


@pytest.mark.parametrize("width", [1, 2, 16])
def test_retries_pure_and_time_reading_policies_against_oracle(env, width):
    """A small cluster under a heavy trace: most events are retries.  Every counter, the placement log and the score
    against the oracle, for policies with and without the template prefix, pure and time-reading, one warp and teams."""
    from helpers import oracle_specs, outcome_digest
    from oracle import pysim
    from funsearch_kubernetes_simulator_b200 import _native as nat
    pols = _retry_policies()
    names = list(pols)
    for cname, n_nodes, tname, head in (("gpu16", 12, "default", 3000), ("gpu1213", 40, "cpu300", 8000), ("all1523", 48, "gpushare100", 5000)):
        cluster = env["ClusterSpec"].load(cname).head(n_nodes)
        tr = env["Trace"].load(tname).head(head)
        ev = env["BatchEvaluator"](cluster, [tr])
        ev.warps_per_team = width
        compiled = ev.compile([pols[n] for n in names])
        assert all(c is not None for c in compiled)
        flags = {n: c.flags for n, c in zip(names, compiled)}
        assert flags["prefixed_pure"] == nat.KS_PF_PREFIX_HOISTED and flags["raw_own_feasibility"] == 0
        assert flags["prefixed_reads_time"] == nat.KS_PF_PREFIX_HOISTED | nat.KS_PF_READS_TIME
        assert flags["raw_reads_time"] == nat.KS_PF_READS_TIME | nat.KS_PF_READS_DURATION
        assert flags["prefixed_reads_duration"] == nat.KS_PF_PREFIX_HOISTED | nat.KS_PF_READS_DURATION
        outs = ev.run_episodes(compiled, list(range(len(names))), [0] * len(names), want_placements=True)
        cs, ts = oracle_specs(cluster, tr)
        retried = 0
        for n, out in zip(names, outs):
            ref = pysim.run_episode(cs, ts, pysim.scorer_from_source(pols[n]))
            tag = (cname, n, width)
            assert (out.status, out.events, out.attempts, out.placed) == (ref.status, ref.events, ref.attempts, ref.placed), tag
            assert (out.max_nodes, out.max_waiting, out.last_event_time) == (ref.max_nodes, ref.max_waiting, ref.last_event_time), tag
            assert outcome_digest(tr, cluster, out) == pysim.placement_digest(ts, cs, ref), tag
            r = out.results
            assert (r.num_snapshots, r.num_fragmentation_events) == (ref.num_snapshots, ref.num_fragmentation_events), tag
            for got, want in ((r.gpu_fragmentation_score, ref.fragmentation), (out.policy_score, ref.score)):
                assert abs(got - want) <= 1e-9 * max(abs(want), 1e-300), (tag, got, want)
            retried += ref.attempts - ref.placed
        assert retried > 1000, "the workload is meant to retry a lot"
