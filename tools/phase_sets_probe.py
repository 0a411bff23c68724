"""Scratch: per-episode phase sums of a KS_PHASE_CLOCKS=<set> build, bench population, 1 and 7 warps per SM."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, build
build.build()
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
names = {"1": ("pop", "score"), "2": ("fetch", "fail path"), "3": ("place path", "update+push+record")}[os.environ["KS_PHASE_CLOCKS"]]
pop = policy_population(1024, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
for n_ep in (148, 1024):
    rb = ev.prepare_resident(compiled, list(range(n_ep)), [0] * n_ep)
    rb.launch(); torch.cuda.synchronize(); rb.launch(); torch.cuda.synchronize()
    r = rb.results()
    ev_ = r["events"].astype(float); at = r["attempts"].astype(float); placed = r["placed"].astype(float)
    a = r["reserved"][:, 0].astype(float); b = r["reserved"][:, 1].astype(float); cyc = r["device_cycles"].astype(float)
    fails = at - placed
    print(f"n={n_ep}: cyc/event {np.mean(cyc/ev_):.0f} | {names[0]}: {np.mean(a/ev_):.0f}/event ({np.mean(a/np.maximum(placed,1)):.0f}/placement) | "
          f"{names[1]}: {np.mean(b/ev_):.0f}/event ({np.mean(b/np.maximum(fails,1)):.0f}/failed attempt) | events {ev_.mean():.0f} attempts {at.mean():.0f} fails {fails.mean():.0f}", flush=True)
    # what a placement, a failed attempt (retry) and a deletion cost: least squares over the episodes
    ok = r["status"] == 0
    X = np.stack([placed[ok], fails[ok], (ev_ - at)[ok]], axis=1)
    for nm, y in (("total", cyc[ok]), (names[0], a[ok]), (names[1], b[ok])):
        coef, *_ = np.linalg.lstsq(X, y, rcond=None)
        print(f"    {nm:>20s}: {coef[0]:7.0f} cycles/placement  {coef[1]:7.0f} /failed attempt  {coef[2]:7.0f} /deletion", flush=True)
