"""Scratch: policy-major table, one vs two CTAs per SM (512 policies x 16 bootstrap traces)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.workload import bootstrap_trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
NP, NT = 512, 16
pop = policy_population(NP, seed=0)
base = Trace.load("default")
traces = [bootstrap_trace(base, s) for s in range(NT)]
ev = BatchEvaluator(ClusterSpec.load("gpu16"), traces)
compiled = ev.compile(pop)
ep_pol = [p for p in range(NP) for t in range(NT)]; ep_tr = [t for p in range(NP) for t in range(NT)]
for ctas in (1, 2):
    ev.ctas_per_sm = ctas
    rb = ev.prepare_resident(compiled, ep_pol, ep_tr)
    best = 1e9
    for it in range(2):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    r = rb.results()
    cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float)
    print(f"ctas/SM {rb.plan.ctas_per_sm} blocks {rb.plan.blocks} teams {rb.plan.teams_per_block}: kernel {best:8.2f} ms  cyc/event {np.mean(cyc/ev_):6.0f}  placements/s {r['placed'].sum()/best*1e3:.3e} ok {int((r['status']==0).sum())}", flush=True)
