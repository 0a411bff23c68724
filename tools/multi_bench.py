"""Scratch: a population x traces batch (512 policies x 16 bootstrap traces, policy-major, planner's choice of kernel); ncu target."""
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
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [bootstrap_trace(base, s) for s in range(NT)])
compiled = ev.compile(pop)
rb = ev.prepare_resident(compiled, [p for p in range(NP) for t in range(NT)], [t for p in range(NP) for t in range(NT)])
for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    r = rb.results()
    print(f"launch {it}: ctas/SM {rb.plan.ctas_per_sm} blocks {rb.plan.blocks}: {e0.elapsed_time(e1):8.2f} ms  placements/s {r['placed'].sum()/e0.elapsed_time(e1)*1e3:.3e} ok {int((r['status']==0).sum())}/{NP*NT}", flush=True)
