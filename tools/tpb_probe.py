"""configs[1]: fewer resident episodes per SM with the table's tail handed out dynamically (stealing), against the single
wave of 7-8 per SM: each episode runs faster with fewer distinct policies on its SM, the SM as a whole does not.
python tools/tpb_probe.py [seed]"""
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
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rb.launch()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    r = rb.results()
    cyc = r["device_cycles"].astype(float)
    print(f"{tag:34s} {min(ms[1:]):7.2f} ms  blocks {rb.plan.blocks} x {rb.plan.teams_per_block}  slowest {cyc.max() / 1e6:.0f} Mcyc mean {cyc.mean() / 1e6:.0f}  "
          f"cyc/event {np.mean(cyc / np.maximum(r['events'], 1)):.0f}", flush=True)
    return r


ev.pack_by_cost = True
base = run("packed by code size, one wave", ev.prepare_resident(compiled, list(range(n)), [0] * n))
ev.pack_by_cost = False
for tpb in (8, 7, 6, 5, 4, 3):
    r = run(f"stealing, {tpb} per SM", ev.prepare_resident(compiled, list(range(n)), [0] * n, teams_per_block=tpb))
    for f in ("status", "events", "attempts", "placed", "policy_score"):
        assert np.array_equal(base[f], r[f]), f
print("results identical")
