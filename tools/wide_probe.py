"""Scratch: where does an event's time go on a 1,024-node cluster (teams of W warps)?  Needs KS_PHASE_CLOCKS=1 for the split."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, build
build.build()
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
pop = policy_population(64, seed=0)
trace = Trace.load("default")
big = ClusterSpec.load("all1523").head(1024)
ev = BatchEvaluator(big, [trace])
compiled = ev.compile(pop)
for W in (8, 16):
    ev.warps_per_team = W
    rb = ev.prepare_resident(compiled, list(range(64)), [0] * 64)
    rb.launch(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    r = rb.results()
    cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float); at = r["attempts"].astype(float)
    pp = r["reserved"][:, 0].astype(float); sc = r["reserved"][:, 1].astype(float)
    print(f"W={W:2d} teams/blk {rb.plan.teams_per_block} blocks {rb.plan.blocks}: {ms:8.2f} ms placements/s {r['placed'].sum()/ms*1e3:.3e} | events {ev_.mean():.0f} attempts {at.mean():.0f} "
          f"cyc/event {np.mean(cyc/ev_):.0f} pop {np.mean(pp/ev_):.0f} score/attempt {np.mean(sc/at):.0f} other/event {np.mean((cyc-pp-sc)/ev_):.0f} ok {int((r['status']==0).sum())}", flush=True)
