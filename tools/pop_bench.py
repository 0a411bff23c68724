"""Scratch: the bench population (1,024 distinct policies x default trace x 16 nodes), a few launches; the target of ncu captures."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
pop = policy_population(n, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
ev.bind(compiled)
ev.wait_background()   # the generation's structures relinked into one module: the steady state bench.py times
rb = ev.prepare_resident(compiled, list(range(n)), [0] * n)
print("modules in batch", rb.n_modules, "structures", rb.n_structures, flush=True)
for it in range(reps):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    r = rb.results()
    cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float)
    print(f"launch {it}: {e0.elapsed_time(e1):7.2f} ms  placements/s {r['placed'].sum()/e0.elapsed_time(e1)*1e3:.3e}  cyc/event {np.mean(cyc/ev_):.0f}  ok {int((r['status']==0).sum())}/{n}", flush=True)
