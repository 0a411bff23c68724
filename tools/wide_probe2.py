"""Scratch: throughput on a 1,024-node cluster with a full population (296 episodes) for several team shapes."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, build
build.build()
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
NP = int(sys.argv[1]) if len(sys.argv) > 1 else 296
pop = policy_population(NP, seed=0)
trace = Trace.load("default")
big = ClusterSpec.load("all1523").head(1024)
ev = BatchEvaluator(big, [trace])
compiled = ev.compile(pop)
for W, T in ((16, 1), (8, 2), (8, 1), (4, 3), (4, 2), (2, 3), (1, 3)):
    ev.warps_per_team = W
    try:
        rb = ev.prepare_resident(compiled, list(range(NP)), [0] * NP, teams_per_block=T)
    except Exception as e:
        print(f"W={W} T={T}: {e}"); continue
    rb.launch(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    r = rb.results()
    cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float)
    print(f"W={W:2d} teams/blk {rb.plan.teams_per_block} blocks {rb.plan.blocks} smem/blk {rb.plan.smem_per_block}: {ms:8.2f} ms placements/s {r['placed'].sum()/ms*1e3:.3e} "
          f"cyc/event {np.mean(cyc/ev_):.0f} ok {int((r['status']==0).sum())}", flush=True)
