"""The sharded product API on real NCCL ranks: distributed.evaluate_population_sharded with BatchEvaluator.evaluate_population
as the local evaluator, two ranks on two GPUs, against the single-GPU result.  Skipped below two visible GPUs."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _worker(rank, world, port, codes, costs, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.distributed import evaluate_population_sharded
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    try:
        ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default").head(1200), Trace.load("cpu100").head(900)],
                            device=rank, processes=2)
        scores = evaluate_population_sharded(codes, ev.evaluate_population, costs, device=torch.device("cuda", rank))
        q.put((rank, scores))
    finally:
        dist.destroy_process_group()


def test_sharded_population_on_two_gpus_matches_one_gpu(built_lib):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    import torch.multiprocessing as mp
    sys.path.insert(0, ROOT)
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    codes = policy_population(41, seed=31) + ["def priority_function(pod, node):\n    return 1 / 0\n"]
    costs = [float(1 + (7 * i) % 13) for i in range(len(codes))]
    costs[3] = 0.0
    ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default").head(1200), Trace.load("cpu100").head(900)], processes=2)
    want = ev.evaluate_population(codes)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29641, codes, costs, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, scores in got:
        assert scores == want, f"rank {rank}"
    assert want[-1] == 0.0 and sum(1 for s in want if s > 0) > 30
