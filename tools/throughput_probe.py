"""Scratch: many-episode (throughput) regime for ncu: 8192 episodes, 64 policies x 128 copies, sorted by policy."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
pop = policy_population(64, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
ev.ctas_per_sm = int(sys.argv[1]) if len(sys.argv) > 1 else 1
n = 8192
ep = sorted(i % 64 for i in range(n))
rb = ev.prepare_resident(compiled, ep, [0] * n, teams_per_block=8)
for it in range(2):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    r = rb.results()
    print(f"ctas/sm={rb.plan.ctas_per_sm} blocks={rb.plan.blocks}: {e0.elapsed_time(e1):.1f} ms placements/s {r['placed'].sum()/e0.elapsed_time(e1)*1e3:.3e}")
