"""Scratch: which episodes of the bench population set the kernel time?"""
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
rb = ev.prepare_resident(compiled, list(range(1024)), [0] * 1024)
rb.launch(); torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
r = rb.results()
ms = r["device_cycles"] / 1.965e6
print(f"kernel {e0.elapsed_time(e1):.1f} ms; episode ms: mean {ms.mean():.1f} p50 {np.median(ms):.1f} p90 {np.percentile(ms,90):.1f} p99 {np.percentile(ms,99):.1f} max {ms.max():.1f}")
print("sum of episode time / (148 SMs x kernel time) = %.2f warps busy per SM on average" % (ms.sum() / (148 * e0.elapsed_time(e1))))
order = np.argsort(-ms)[:12]
for i in order:
    print(f"ep {i:4d}: {ms[i]:6.1f} ms events {r['events'][i]:6d} attempts {r['attempts'][i]:6d} frag {r['num_fragmentation_events'][i]:6d} cyc/event {r['device_cycles'][i]/r['events'][i]:6.0f} score {r['policy_score'][i]:.4f}")
    body = pop[i].split("score = 0.0")[1] if "score = 0.0" in pop[i] else pop[i]
    print("      " + " | ".join(l.strip() for l in body.strip().splitlines()[:4])[:200])
