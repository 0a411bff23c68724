"""Scratch: per-phase cycle breakdown (needs a build with KS_PHASE_CLOCKS=1)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, build
build.build()
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
pol = json.load(open(os.path.join(ROOT, "tests/golden/policies.json")))
names = [n for n in sorted(pol) if not n.startswith(("x_", "sim_"))]
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile([pol[n] for n in names])
for n_ep in (148, 1024):
    rb = ev.prepare_resident(compiled, [i % len(names) for i in range(n_ep)], [0] * n_ep)
    rb.launch(); r = rb.results()
    print(f"--- {n_ep} episodes")
    for k, n in enumerate(names):
        sel = [i for i in range(n_ep) if i % len(names) == k]
        ev_ = r["events"][sel].astype(float); at = r["attempts"][sel].astype(float)
        cyc = r["device_cycles"][sel] / ev_; pop = r["reserved"][sel, 0] / ev_; sc = r["reserved"][sel, 1] / at
        print(f"{n:24s} cyc/event {cyc.mean():7.0f} | pop/event {pop.mean():6.0f} | score/attempt {sc.mean():7.0f} | attempts/event {np.mean(at/ev_):.2f} | other/event {np.mean(cyc - pop - sc*at/ev_):6.0f}")
