"""configs[1]: the default cut of the episode table against per-SM ranges packed by predicted cost (proxy: code size; or
the cycles an earlier replay measured).  python tools/pack_probe.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace  # noqa: E402
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator  # noqa: E402
from funsearch_kubernetes_simulator_b200.variants import policy_population  # noqa: E402

n = 1024
pop = policy_population(n, seed=int(sys.argv[1]) if len(sys.argv) > 1 else 0)
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
compiled = ev.compile(pop)
ev.bind(compiled)
ev.wait_background()


def run(tag, rb):
    ms = []
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rb.launch()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    r = rb.results()
    cyc = r["device_cycles"].astype(float)
    print(f"{tag:28s} {min(ms[1:]):7.2f} ms  blocks {rb.plan.blocks} x {rb.plan.teams_per_block}  slowest {cyc.max() / 1e6:.0f} Mcyc mean {cyc.mean() / 1e6:.0f}"
          f"  predicted makespan/heaviest {getattr(rb.tables, 'predicted_makespan_over_heaviest', 0):.2f}", flush=True)
    return r


ev.pack_by_cost = False
base = run("default cut", ev.prepare_resident(compiled, list(range(n)), [0] * n))
ev.pack_by_cost = True
r1 = run("packed by code size", ev.prepare_resident(compiled, list(range(n)), [0] * n))
r2 = run("packed by measured cycles", ev.prepare_resident(compiled, list(range(n)), [0] * n, costs=base["device_cycles"].astype(float)))
r3 = run("packed by measured events", ev.prepare_resident(compiled, list(range(n)), [0] * n, costs=base["events"].astype(float)))
for f in ("status", "events", "attempts", "placed", "policy_score"):
    assert np.array_equal(base[f], r1[f]) and np.array_equal(base[f], r2[f]) and np.array_equal(base[f], r3[f]), f
print("results identical")
