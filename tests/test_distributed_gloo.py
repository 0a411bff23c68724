"""The N>1 path on CPU: world_size-2 gloo processes shard a population, score their shard (with the oracle on a
tiny workload standing in for the GPU evaluator) and all-gather the fitness; every rank must end up with exactly
the single-process result."""
import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _oracle_scores(codes):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_specs
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from oracle import pysim
    cs, ts = oracle_specs(ClusterSpec.load("gpu16"), Trace.load("default").head(400))
    return [pysim.run_episode(cs, ts, pysim.scorer_from_source(c)).score for c in codes]


def _worker(rank, world, port, codes, costs, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from funsearch_kubernetes_simulator_b200.distributed import evaluate_population_sharded, episode_shard
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        scores = evaluate_population_sharded(codes, _oracle_scores, costs)
        q.put((rank, scores, episode_shard(len(codes), rank, world, costs)))
    finally:
        dist.destroy_process_group()


COSTS = {"none": None, "mixed": [5.0, 1.0, 9.0, 2.0, 2.0, 7.0, 1.0],
         # shapes that once made shards larger than the gather block (and one rank raise before the collective):
         "zeros": [0.0, 0.0, 0.0, 0.0, 3.0, 0.0, 0.0],      # last generation's failed policies cost nothing
         "dominant": [1e6, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0]}   # one episode outweighs all others together


@pytest.mark.parametrize("with_costs", ["none", "mixed", "zeros", "dominant"])
def test_sharded_population_matches_single_process(with_costs):
    import torch.multiprocessing as mp
    sys.path.insert(0, ROOT)
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    codes = policy_population(7, seed=2)          # odd count: ragged shards
    costs = COSTS[with_costs]
    want = _oracle_scores(codes)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29611 + list(COSTS).index(with_costs)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, codes, costs, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    shards = {}
    for rank, scores, shard in got:
        assert scores == want, f"rank {rank}"
        shards[rank] = shard
    assert sorted(shards[0] + shards[1]) == list(range(7)) and not set(shards[0]) & set(shards[1])
    if with_costs == "mixed":   # LPT keeps the two ranks' loads close
        load = [sum(costs[i] for i in shards[r]) for r in (0, 1)]
        assert abs(load[0] - load[1]) <= 3.0


def test_shard_helpers_single_process():
    from funsearch_kubernetes_simulator_b200.distributed import episode_shard, gather_scores, shard_plan
    assert episode_shard(10, 1, 4) == [1, 5, 9]
    assert episode_shard(0, 0, 2) == []
    all_idx = sorted(sum((episode_shard(11, r, 3, list(range(11))) for r in range(3)), []))
    assert all_idx == list(range(11))
    assert gather_scores([0.5, 0.25], [1, 0], 2) == [0.25, 0.5]
    # zero-cost items spread over the ranks instead of piling onto the lightest one
    plan = shard_plan(64, 4, [0.0] * 30 + [float(i + 1) for i in range(34)])
    assert sorted(sum(plan, [])) == list(range(64)) and max(len(sh) for sh in plan) - min(len(sh) for sh in plan) <= 2
    # one dominant item: it gets a rank to itself, the others share the rest evenly
    plan = shard_plan(64, 4, [1e6] + [1.0] * 63)
    assert sorted(len(sh) for sh in plan) == [1, 21, 21, 21]
