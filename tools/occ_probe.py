"""Scratch: the bench population at several occupancies, plus same-policy controls (for ncu metric passes)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
pop = policy_population(1024, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
def run(tag, sel):
    rb = ev.prepare_resident(compiled, sel, [0] * len(sel))
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    r = rb.results()
    cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float)
    print(f"{tag:40s} n={len(sel):5d} kernel {e0.elapsed_time(e1):7.1f} ms  cyc/event mean {np.mean(cyc/ev_):6.0f}  Mcyc mean {cyc.mean()/1e6:6.1f} max {cyc.max()/1e6:6.1f}", flush=True)
run("warm", list(range(148)))
run("population, 148", list(range(148)))
run("population, 592", list(range(592)))
run("population, 1024", list(range(1024)))
run("policy 707 x1024 (heavy pow)", [707] * 1024)
run("policy 707 x148", [707] * 148)
run("policy 5 x1024 (light)", [5] * 1024)
run("policy 5 x148", [5] * 148)
run("8 policies cycled x1024", [i % 8 for i in range(1024)])
run("64 policies cycled x1024", [i % 64 for i in range(1024)])
