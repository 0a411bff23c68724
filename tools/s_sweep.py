"""Scratch: sweep the shared-memory share of the event heap (the rest of the 228 KB stays L1D)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
cluster = ClusterSpec.load("gpu16"); trace = Trace.load("default")
ev = BatchEvaluator(cluster, [trace])
pop = policy_population(1024, seed=0)
compiled = ev.compile(pop)
ep = list(range(1024))
for teams, S in [(7, 0), (7, 2), (7, 64), (7, 256), (7, 1024), (7, 2048), (7, 3904), (8, 64), (8, 1024), (4, 64)]:
    rb = ev.prepare_resident(compiled, ep, [0] * 1024, teams_per_block=teams, heap_smem_entries=S)
    import torch
    ts = []
    for it in range(3):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    r = rb.results()
    print(f"teams={rb.plan.teams_per_block} blocks={rb.plan.blocks} S={rb.plan.heap_smem_entries:5d} smem/block={rb.plan.smem_per_block:6d}  ms {min(ts):7.2f}  placements/s {r['placed'].sum()/min(ts)*1e3:.3e} ok={int((r['status']==0).sum())}")
