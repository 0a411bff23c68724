"""Which episodes of the bench population are slow, and why: top episodes by device cycles with their counters and source;
episodes whose status differs from the oracle's are printed too (python tools/slow_probe.py [seed] [n])."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace  # noqa: E402
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator  # noqa: E402
from funsearch_kubernetes_simulator_b200.variants import policy_population  # noqa: E402

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 0
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
pop = policy_population(n, seed=seed)
cluster, trace = ClusterSpec.load("gpu16"), Trace.load("default")
ev = BatchEvaluator(cluster, [trace])
compiled = ev.compile(pop)
rb = ev.prepare_resident(compiled, list(range(n)), [0] * n)
rb.launch()
res = rb.results()
cyc = res["device_cycles"].astype(np.float64)
order = np.argsort(-cyc)
print("mean Mcycles", cyc.mean() / 1e6, "median", np.median(cyc) / 1e6, "p99", np.percentile(cyc, 99) / 1e6)
print("status histogram", {int(s): int((res["status"] == s).sum()) for s in np.unique(res["status"])})
for i in order[:6]:
    r = res[i]
    print(f"--- episode {i}: {cyc[i] / 1e6:.0f} Mcycles, status {r['status']}, events {r['events']}, attempts {r['attempts']}, "
          f"placed {r['placed']}, frag events {r['num_fragmentation_events']}, max_wait {r['max_waiting']}, "
          f"cycles/attempt {cyc[i] / max(1, r['attempts']):.0f}")
    print(pop[i])
bad = [int(i) for i in np.nonzero(res["status"] != 0)[0]]
print("failed episodes:", bad[:40])
for i in bad[:8]:
    print(f"--- failed {i}: status {res[i]['status']} events {res[i]['events']} attempts {res[i]['attempts']}")
    print(pop[i])
