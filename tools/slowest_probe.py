"""Scratch: the slowest episodes of the bench population (needs KS_PHASE_CLOCKS=1)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, build
build.build()
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
n = 1024
pop = policy_population(n, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
rb = ev.prepare_resident(compiled, list(range(n)), [0] * n)
rb.launch(); torch.cuda.synchronize(); rb.launch(); torch.cuda.synchronize()
r = rb.results()
cyc = r["device_cycles"].astype(float); ev_ = r["events"].astype(float); at = r["attempts"].astype(float)
pp = r["reserved"][:, 0].astype(float); sc = r["reserved"][:, 1].astype(float)
print("corr(total, events) %.2f corr(total, score/attempt) %.2f" % (np.corrcoef(cyc, ev_)[0, 1], np.corrcoef(cyc, sc / at)[0, 1]))
print("events: min %d p50 %d p90 %d max %d" % (ev_.min(), np.median(ev_), np.percentile(ev_, 90), ev_.max()))
for i in np.argsort(-cyc)[:12]:
    src = pop[i].split("score = 0.0")[-1]
    print(f"ep {i:4d}: {cyc[i]/1e6:6.1f} Mcyc events {ev_[i]:6.0f} attempts {at[i]:6.0f} pop/ev {pp[i]/ev_[i]:5.0f} score/att {sc[i]/at[i]:6.0f} other/ev {(cyc[i]-pp[i]-sc[i])/ev_[i]:5.0f} | pow {src.count('**')} div {src.count('/')} sorted {src.count('sorted')} for {src.count(' for ')} sqrt {src.count('sqrt')}")
i = int(np.argsort(-cyc)[0]); print(pop[i])
