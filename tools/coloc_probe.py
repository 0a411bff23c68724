"""Scratch: same-policy episodes co-located on an SM vs spread (148 policies x 8 bootstrap traces, one wave)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.workload import bootstrap_trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
NP, NT = 148, 8
pop = policy_population(NP, seed=0)
base = Trace.load("default")
traces = [bootstrap_trace(base, s) for s in range(NT)]
ev = BatchEvaluator(ClusterSpec.load("gpu16"), traces)
compiled = ev.compile(pop)
def run(tag, ep_pol, ep_tr):
    rb = ev.prepare_resident(compiled, ep_pol, ep_tr, teams_per_block=8)
    best = 1e9
    for it in range(3):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    r = rb.results()
    cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float)
    print(f"{tag:40s} blocks {rb.plan.blocks}: kernel {best:7.2f} ms  cyc/event {np.mean(cyc/ev_):6.0f}  placements/s {r['placed'].sum()/best*1e3:.3e} ok {int((r['status']==0).sum())}", flush=True)
# policy-major: block b = policy b on 8 traces
run("same policy per SM", [p for p in range(NP) for t in range(NT)], [t for p in range(NP) for t in range(NT)])
# trace-major: block b gets 8 different policies
run("8 different policies per SM", [p for t in range(NT) for p in range(NP)], [t for t in range(NT) for p in range(NP)])
# multi-wave: 512 policies x 8 traces = 4,096 episodes
NP2 = 512
pop2 = policy_population(NP2, seed=0)
compiled = ev.compile(pop2)
run("4096 ep, policy-major table", [p for p in range(NP2) for t in range(NT)], [t for p in range(NP2) for t in range(NT)])
run("4096 ep, trace-major table", [p for t in range(NT) for p in range(NP2)], [t for t in range(NT) for p in range(NP2)])
