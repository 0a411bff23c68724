"""BASELINE configs[2] shape: 64 policies x a synthetic Alibaba-shaped trace x 1,024 nodes, by team width
(python tools/config2_probe.py [pods] [widths...]; KS_EVENT_BUDGET=<n> caps the events of an episode like bench.py --config 2;
with a KS_PHASE_CLOCKS build (KS_B200_LIB=<that .so>) the phase sums of the slowest episodes are printed)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from funsearch_kubernetes_simulator_b200 import ClusterSpec, synthetic_trace  # noqa: E402
from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator  # noqa: E402
from funsearch_kubernetes_simulator_b200.variants import policy_population  # noqa: E402

pods = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
widths = [int(x) for x in sys.argv[2:]] or [16, 8, 4]
pop = policy_population(64, seed=0)
trace = synthetic_trace(pods, seed=3, load_factor=200.0)
big = ClusterSpec.load("all1523").head(1024)
ev = BatchEvaluator(big, [trace])
compiled = ev.compile(pop)
assert all(c is not None for c in compiled)
for W in widths:
    ev.warps_per_team = W
    rb = ev.prepare_resident(compiled, list(range(64)), [0] * 64, single_module=True, max_events=int(os.environ.get("KS_EVENT_BUDGET", "0")))
    rb.launch()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    rb.launch()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    r = rb.results()
    cyc = r["device_cycles"].astype(float)
    ev_ = np.maximum(r["events"].astype(float), 1)
    pa, pb = r["reserved"][:, 0].astype(float), r["reserved"][:, 1].astype(float)
    if pa.sum() or pb.sum():
        at = np.maximum(r["attempts"].astype(float), 1)
        print(f"   phase clocks (KS_PHASE_CLOCKS={os.environ.get('KS_PHASE_CLOCKS')}): a/event {np.mean(pa / ev_):.0f}  b/event {np.mean(pb / ev_):.0f}  "
              f"a/attempt {np.mean(pa / at):.0f}  b/attempt {np.mean(pb / at):.0f}  max_wait mean {r['max_waiting'].mean():.0f} "
              f"frag events/ep {r['num_fragmentation_events'].mean():.0f}")
    okk = np.nonzero(r['status'] == 0)[0]
    for i in list(np.argsort(-cyc)[:3]) + list(okk[np.argsort(-cyc[okk])][:2]):
        print(f"   slow ep {i} status {r['status'][i]}: {cyc[i] / 1e6:.0f} Mcyc events {r['events'][i]} attempts {r['attempts'][i]} placed {r['placed'][i]} frag {r['num_fragmentation_events'][i]} "
              f"max_wait {r['max_waiting'][i]} a/event {pa[i] / ev_[i]:.0f} b/event {pb[i] / ev_[i]:.0f} "
              f"cyc/event {cyc[i] / ev_[i]:.0f} hoisted {compiled[i].flags} consts {len(compiled[i].consts)} src lines {pop[i].count(chr(10))}")
    print(f"   all episodes: sum(cyc)/sum(events) = {cyc.sum() / ev_.sum():.0f}")
    print(f"W={W:2d} teams/blk {rb.plan.teams_per_block} blocks {rb.plan.blocks} smem/blk {rb.plan.smem_per_block}: {ms:9.1f} ms  "
          f"placements/s {r['placed'].sum() / ms * 1e3:.3e} | events/ep {ev_.mean():.0f} attempts/ep {r['attempts'].mean():.0f} "
          f"cyc/event mean {np.mean(cyc / ev_):.0f} slowest ep {cyc.max() / 1e6:.0f} Mcyc  status ok {int((r['status'] == 0).sum())}/64",
          flush=True)
