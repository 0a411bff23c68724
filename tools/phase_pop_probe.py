"""Scratch: per-phase cycle breakdown over the bench population at several occupancies (needs KS_PHASE_CLOCKS=1)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, build
build.build()
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
pop = policy_population(1024, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
for n_ep in (148, 296, 592, 1024):
    sel = list(range(n_ep))
    rb = ev.prepare_resident(compiled, sel, [0] * n_ep)
    rb.launch(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    r = rb.results()
    ev_ = r["events"].astype(float); at = r["attempts"].astype(float)
    cyc = r["device_cycles"].astype(float); pp = r["reserved"][:, 0].astype(float); sc = r["reserved"][:, 1].astype(float)
    print(f"--- {n_ep} episodes: kernel {e0.elapsed_time(e1):.1f} ms; episode Mcyc mean {cyc.mean()/1e6:.1f} max {cyc.max()/1e6:.1f}; "
          f"events {ev_.mean():.0f} attempts {at.mean():.0f}; per event: total {np.mean(cyc/ev_):.0f} pop {np.mean(pp/ev_):.0f} "
          f"other {np.mean((cyc-pp-sc)/ev_):.0f}; score/attempt mean {np.mean(sc/at):.0f} p50 {np.median(sc/at):.0f} p90 {np.percentile(sc/at,90):.0f} max {np.max(sc/at):.0f}")
    if n_ep == 1024:
        o = np.argsort(-(sc / at))[:5]
        for i in o:
            print(f"   ep {i}: score/attempt {sc[i]/at[i]:.0f} attempts {at[i]:.0f} events {ev_[i]:.0f} total Mcyc {cyc[i]/1e6:.1f}")
            print("      " + " | ".join(l.strip() for l in pop[i].split("score = 0.0")[-1].strip().splitlines()[:8])[:400])
