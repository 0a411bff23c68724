"""Scratch timing of the replay kernel (not the contract bench): N episodes cycling the golden policies."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator

n_ep = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
teams = int(sys.argv[2]) if len(sys.argv) > 2 else 0
S = int(sys.argv[3]) if len(sys.argv) > 3 else 0
pol = json.load(open(os.path.join(ROOT, "tests/golden/policies.json")))
names = [n for n in sorted(pol) if not n.startswith(("x_", "sim_"))]
cluster = ClusterSpec.load("gpu16"); trace = Trace.load("default")
ev = BatchEvaluator(cluster, [trace])
t0 = time.time(); compiled = ev.compile([pol[n] for n in names]); print("compile s", time.time() - t0)
ep_pol = [i % len(names) for i in range(n_ep)]
for it in range(3):
    t0 = time.time()
    outs = ev.run_episodes(compiled, ep_pol, [0] * n_ep, teams_per_block=teams)
    wall = time.time() - t0
    placed = sum(o.placed for o in outs); evals = sum(o.scorer_calls for o in outs); events = sum(o.events for o in outs)
    ms = ev.last_kernel_ms
    p = ev.last_plan
    print(f"iter {it}: kernel {ms:.2f} ms wall {wall*1e3:.1f} ms | placements/s {placed/ms*1e3:.3e} evals/s {evals/ms*1e3:.3e} "
          f"events/s {events/ms*1e3:.3e} | blocks {p.blocks} teams {p.teams_per_block} smem/block {p.smem_per_block}")
cyc = np.array([o.device_cycles for o in outs]); evs = np.array([o.events for o in outs])
print("cycles/event: mean %.0f min %.0f max %.0f" % ((cyc/evs).mean(), (cyc/evs).min(), (cyc/evs).max()))
for n in names:
    idx = [i for i in range(n_ep) if names[i % len(names)] == n][:1]
    o = outs[idx[0]]; print(f"  {n:28s} st={o.status} events={o.events} cyc/event={o.device_cycles/max(o.events,1):.0f} score={o.policy_score:.6f}")
