#!/usr/bin/env python3
"""BASELINE configs 4 and 5 on ONE GPU (the per-GPU share; ranks are independent, see bench.py --gpus N):
  config 5: episode-count sweep 2^10 .. 2^17 (policy, trace) episodes of the config-2 shape
  config 4: 512 policies x 16 bundled traces x 4 seeds = 32,768 episodes (1/8 of the 262,144-episode population)
Writes one JSON object per line to stdout; `python tools/sweep.py > profiles/sweep_r01.jsonl`."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, available_traces, bootstrap_trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population


def timed(rb, reps=1):
    best = None
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return best


cluster = ClusterSpec.load("gpu16")
pop = policy_population(1024, seed=0)
# ---- config 5
ev = BatchEvaluator(cluster, [Trace.load("default")])
compiled = ev.compile(pop)
for logn in range(10, 18):
    n = 1 << logn
    ep = sorted(i % 1024 for i in range(n))
    rb = ev.prepare_resident(compiled, ep, [0] * n)
    rb.launch(); torch.cuda.synchronize()
    ms = timed(rb)
    r = rb.results()
    print(json.dumps({"config": 5, "episodes": n, "ms": ms, "placements_per_s": float(r["placed"].sum()) / ms * 1e3,
                      "policy_evals_per_s": float(r["scorer_calls"].sum()) / ms * 1e3, "episodes_per_s": n / ms * 1e3,
                      "ok": int((r["status"] == 0).sum()), "teams_per_block": rb.plan.teams_per_block, "blocks": rb.plan.blocks, "ctas_per_sm": rb.plan.ctas_per_sm}), flush=True)
# ---- config 4 (one GPU's share)
names = [t for t in available_traces() if t not in ("cpu037", "cpu072", "gpuspec05")][:16]
traces = [bootstrap_trace(Trace.load(t), s) for t in names for s in range(4)]
ev4 = BatchEvaluator(cluster, traces)
comp4 = ev4.compile(pop[:512])
ep_pol = [p for p in range(512) for _ in range(len(traces))]
ep_tr = [t for _ in range(512) for t in range(len(traces))]
rb = ev4.prepare_resident(comp4, ep_pol, ep_tr)
rb.launch(); torch.cuda.synchronize()
ms = timed(rb)
r = rb.results()
fit = r["policy_score"].reshape(512, len(traces)).mean(axis=1)
print(json.dumps({"config": 4, "episodes": len(ep_pol), "policies": 512, "traces": len(names), "seeds": 4, "ms": ms,
                  "placements_per_s": float(r["placed"].sum()) / ms * 1e3, "episodes_per_s": len(ep_pol) / ms * 1e3,
                  "status_counts": {str(k): int((r["status"] == k).sum()) for k in range(6)},
                  "ctas_per_sm": rb.plan.ctas_per_sm, "best_fitness": float(fit.max()), "mean_fitness": float(fit.mean())}), flush=True)
