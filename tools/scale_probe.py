"""Scratch: (1) team width on a 1,024-node cluster, (2) throughput vs number of episodes on the 16-node cluster."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
def timeit(rb, n=2):
    ts = []
    for it in range(n):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts)
pop = policy_population(1024, seed=0)
trace = Trace.load("default")
which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "wide"):
    big = ClusterSpec.load("all1523").head(1024)
    ev = BatchEvaluator(big, [trace])
    compiled = ev.compile(pop[:64])
    for W in (1, 4, 8, 16):
        ev.warps_per_team = W
        rb = ev.prepare_resident(compiled, list(range(64)), [0] * 64)
        ms = timeit(rb)
        r = rb.results()
        print(f"N=1024 W={W:2d} teams/blk={rb.plan.teams_per_block} blocks={rb.plan.blocks}: {ms:9.2f} ms  placements/s {r['placed'].sum()/ms*1e3:.3e}  evals/s {r['scorer_calls'].sum()/ms*1e3:.3e} ok={int((r['status']==0).sum())}")
if which in ("all", "many"):
    ev = BatchEvaluator(ClusterSpec.load("gpu16"), [trace])
    compiled = ev.compile(pop)
    for n_ep, teams, ctas in ((1024, 0, 1), (1024, 0, 2), (4096, 8, 1), (4096, 8, 2), (16384, 8, 1), (16384, 8, 2), (16384, 4, 2)):
        ev.ctas_per_sm = ctas
        ep = [i % 1024 for i in range(n_ep)]
        # episodes of one policy adjacent: sort by policy so that co-resident warps share code
        ep.sort()
        rb = ev.prepare_resident(compiled, ep, [0] * n_ep, teams_per_block=teams)
        ms = timeit(rb, 1)
        r = rb.results()
        print(f"n_ep={n_ep:6d} ctas/sm={rb.plan.ctas_per_sm} teams/blk={rb.plan.teams_per_block} blocks={rb.plan.blocks}: {ms:9.2f} ms  placements/s {r['placed'].sum()/ms*1e3:.3e} ok={int((r['status']==0).sum())}")
