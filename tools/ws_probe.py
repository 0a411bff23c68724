"""Scratch: per-event latency vs resident warps per SM (k teams per block, 148*k distinct policies)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
pop = policy_population(1184, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
for k in (1, 2, 3, 4, 5, 6, 7, 8):
    sel = list(range(148 * k))
    rb = ev.prepare_resident(compiled, sel, [0] * len(sel), teams_per_block=k)
    rb.launch(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    r = rb.results()
    cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float)
    print(f"k={k} warps/SM: n={len(sel):5d} kernel {e0.elapsed_time(e1):7.1f} ms  cyc/event mean {np.mean(cyc/ev_):6.0f}  Mcyc mean {cyc.mean()/1e6:6.1f} max {cyc.max()/1e6:6.1f}", flush=True)
