"""Scratch: is the 16-warps/SM (dense) variant limited by instruction-cache pressure from distinct policies?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
pop = policy_population(1024, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
def run(tag, ep, ctas, teams=8):
    ev.ctas_per_sm = ctas
    rb = ev.prepare_resident(compiled, ep, [0] * len(ep), teams_per_block=teams)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    rb.launch(); torch.cuda.synchronize()
    e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1); r = rb.results()
    print(f"{tag:40s} ctas/sm={ctas} blocks={rb.plan.blocks}: {ms:8.1f} ms  placements/s {r['placed'].sum()/ms*1e3:.3e}")
n = 8192
for ctas in (1, 2):
    run("8192 ep, ONE policy", [3] * n, ctas)
    run("8192 ep, 8 policies", [i % 8 for i in range(n)], ctas)
    run("8192 ep, 1024 policies interleaved", [i % 1024 for i in range(n)], ctas)
    run("8192 ep, 1024 policies sorted", sorted(i % 1024 for i in range(n)), ctas)
