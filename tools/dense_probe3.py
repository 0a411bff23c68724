"""Scratch: 4,096 distinct policies x 1 trace, one vs two CTAs per SM."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
NP = 4096
pop = policy_population(NP, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
for ctas in (1, 2):
    ev.ctas_per_sm = ctas
    rb = ev.prepare_resident(compiled, list(range(NP)), [0] * NP)
    best = 1e9
    for it in range(2):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    r = rb.results()
    cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float)
    print(f"ctas/SM {rb.plan.ctas_per_sm} blocks {rb.plan.blocks} teams {rb.plan.teams_per_block}: kernel {best:8.2f} ms  cyc/event {np.mean(cyc/ev_):6.0f}  placements/s {r['placed'].sum()/best*1e3:.3e} ok {int((r['status']==0).sum())}", flush=True)
