"""Scratch experiment: how much of the per-event latency is instruction-cache thrash between warps that run
different policy functions?  Same episode count, 1 / 2 / 13 distinct policies, and fewer warps per SM."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
pol = json.load(open(os.path.join(ROOT, "tests/golden/policies.json")))
names = [n for n in sorted(pol) if not n.startswith(("x_", "sim_"))]
cluster = ClusterSpec.load("gpu16"); trace = Trace.load("default")
ev = BatchEvaluator(cluster, [trace])
compiled = ev.compile([pol[n] for n in names])
def run(tag, ep_pol, teams=0, S=0):
    n_ep = len(ep_pol)
    for it in range(2):
        outs = ev.run_episodes(compiled, ep_pol, [0] * n_ep, teams_per_block=teams)
    cyc = np.array([o.device_cycles for o in outs], dtype=float); evs = np.array([o.events for o in outs], dtype=float)
    p = ev.last_plan
    print(f"{tag:46s} n_ep={n_ep:5d} teams={p.teams_per_block} blocks={p.blocks} kernel {ev.last_kernel_ms:8.2f} ms  cyc/event mean {np.mean(cyc/evs):7.0f}  placements/s {sum(o.placed for o in outs)/ev.last_kernel_ms*1e3:.3e}")
i4816 = names.index("ts_funsearch_4816"); iff = names.index("ts_first_fit"); i4901 = names.index("ts_funsearch_4901")
run("13 policies cycled, 1024 ep", [i % len(names) for i in range(1024)])
run("all 4816, 1024 ep", [i4816] * 1024)
run("all first_fit, 1024 ep", [iff] * 1024)
run("all 4901, 1024 ep", [i4901] * 1024)
run("4816/4901 alternating", [(i4816, i4901)[i % 2] for i in range(1024)])
run("13 policies, 148 ep (1 warp/SM)", [i % len(names) for i in range(148)])
run("all 4816, 148 ep (1 warp/SM)", [i4816] * 148)
run("13 policies, 296 ep (2 warps/SM)", [i % len(names) for i in range(296)])
run("13 policies, 592 ep (4 warps/SM)", [i % len(names) for i in range(592)])
run("13 policies, 1024 ep, teams=4 (2 waves)", [i % len(names) for i in range(1024)], teams=4)
run("13 policies, 1024 ep, S=8152 teams=3", [i % len(names) for i in range(1024)], teams=3, S=8152)
run("13 policies, 1024 ep, S=64", [i % len(names) for i in range(1024)], S=64)
