"""Scratch: how much does contention-aware placement of episodes on SMs buy?  Oracle costs = measured cycles of a first run."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
from funsearch_kubernetes_simulator_b200.variants import policy_population
n = 1024
pop = policy_population(n, seed=0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
def run(tag, ep_pol, teams):
    ep_pol = np.asarray(ep_pol, dtype=np.int32)
    rb = ev.prepare_resident(compiled, ep_pol, [0] * len(ep_pol), teams_per_block=teams)
    best = 1e9
    for it in range(3):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); rb.launch(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    r = rb.results()
    real = ep_pol >= 0
    cyc = r["device_cycles"][real].astype(float)
    print(f"{tag:44s} blocks {rb.plan.blocks} teams {rb.plan.teams_per_block}: kernel {best:7.2f} ms  episode Mcyc mean {cyc.mean()/1e6:6.1f} max {cyc.max()/1e6:6.1f}  ok {int((r['status'][real]==0).sum())}", flush=True)
    out = np.zeros(n); out[ep_pol[real]] = cyc
    return out
cost = run("table order, planner default", list(range(n)), 0)
run("table order, 8 teams/block", list(range(n)), 8)
def lpt(cost, B, T, power=1.0):
    order = np.argsort(-cost)
    load = np.zeros(B); cnt = np.zeros(B, dtype=int); slots = [[] for _ in range(B)]
    for e in order:
        ok = np.nonzero(cnt < T)[0]
        b = ok[np.argmin(load[ok])]
        slots[b].append(int(e)); load[b] += cost[e] ** power; cnt[b] += 1
    table = []
    for b in range(B):
        table += slots[b] + [-1] * (T - len(slots[b]))
    return table, cnt
for power in (1.0, 2.0, 3.0):
    table, cnt = lpt(cost, 148, 8, power)
    print("teams per block histogram:", np.bincount(cnt))
    run(f"LPT on measured cycles^{power}", table, 8)
