"""Where a fresh generation's wall time goes: evaluate_population on never-seen sources, stage by stage
(python tools/fresh_probe.py [generations])."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, compile_pool  # noqa: E402
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator  # noqa: E402
from funsearch_kubernetes_simulator_b200.variants import policy_population  # noqa: E402

gens = int(sys.argv[1]) if len(sys.argv) > 1 else 3
ev = BatchEvaluator(ClusterSpec.load("gpu16"), [Trace.load("default")])
print("pool", compile_pool.wait_ready())
for k in range(gens):
    pop = policy_population(1024, seed=100003 + 7919 * (k + 1))
    t0 = time.perf_counter()
    compiled = ev.compile(pop)
    t1 = time.perf_counter()
    ok = [c for c in compiled if c is not None]
    rb = ev.prepare_resident(ok, list(range(len(ok))), [0] * len(ok))
    t2 = time.perf_counter()
    rb.launch()
    res = rb.results()
    t3 = time.perf_counter()
    cyc = res["device_cycles"].astype(np.float64)
    worst = int(np.argmax(cyc))
    print(f"gen {k}: compile {t1 - t0:.3f} s  tables+bind {t2 - t1:.3f} s  launch+results {t3 - t2:.3f} s  modules {rb.n_modules}  "
          f"slowest {cyc.max() / 1e6:.0f} Mcyc (events {res['events'][worst]}, status {res['status'][worst]}) mean {cyc.mean() / 1e6:.0f}")
    print("   stats", {k2: (round(v, 3) if isinstance(v, float) else v) for k2, v in ev.last_compile_stats.items()})
    if cyc.max() > 1e9:
        print(ok[worst].source)
    t0 = time.perf_counter()
    ev.evaluate_population(pop)
    print(f"   evaluate_population again (everything cached): {time.perf_counter() - t0:.3f} s, modules loaded {len(ev._modules)}")
    if len(sys.argv) > 2:
        time.sleep(float(sys.argv[2]))   # the caller's think time between generations (the background relink uses it)
