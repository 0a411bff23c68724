#!/usr/bin/env python3
"""Headline benchmark: pod placements/s of one FunSearch generation replayed on B200(s).

Workload (BASELINE.json configs[1]): 1,024 policy variants x the bundled 8,152-pod OpenB trace on the
16-node GPU cluster, one (policy, trace) episode each, per GPU ("scaling": "weak": every rank evaluates
its own 1,024-policy population shard).  One "step" = one launch of the replay kernel over the whole
batch (+ the NCCL all-gather of per-episode fitness when N > 1).

  python bench.py [--gpus N] [--steps K] [--warmup W]                 # this framework (CUDA)
  python bench.py --impl reference [--steps K] [--warmup W]           # CPU stand-in of the reference
  torchrun ... bench.py --gpus N ...                                   # one rank per GPU

Prints ONE JSON line (rank 0).  `value` = placements/s with inputs resident in HBM (CUDA events on the
launch stream); `e2e` = the same metric through the host-buffer C-ABI call ks_eval_batch_host with the
workload copied host->device and results device->host inside the timed region; `roofline` relates the
algorithmic bytes (32 B per placement) to the measured HBM peak; `cpu_baseline` is the
oracle port (oracle/pysim.py: same algorithm, CPython + heapq like the reference) on the host cores.
NVRTC compilation of the population is reported separately under `compile` (not part of a step).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pod_placements_per_sec"
UNIT = "placements/s"
N_EPISODES = 1024
WORKLOAD = "1024 policy variants x default OpenB trace (8152 pods) x 16 nodes, 1 episode each (configs[1])"


# --------------------------------------------------------------------------------------------- CPU side
_W = {}


def _cpu_init():
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_specs
    _W["specs"] = oracle_specs(ClusterSpec.load("gpu16"), Trace.load("default"))


def _cpu_episode(src):
    from oracle import pysim
    cs, ts = _W["specs"]
    t0 = time.perf_counter()
    r = pysim.run_episode(cs, ts, pysim.scorer_from_source(src))
    return r.placed, r.scorer_calls, r.events, time.perf_counter() - t0


def cpu_run(sources, workers):
    """Replays `sources` on the default workload with the oracle port on `workers` processes.
    Returns (placements, scorer_calls, wall seconds)."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    with ctx.Pool(workers, initializer=_cpu_init) as pool:
        pool.map(_cpu_episode, sources[:workers])  # warm the workers (imports, trace load)
        t0 = time.perf_counter()
        out = pool.map(_cpu_episode, sources, chunksize=1)
        wall = time.perf_counter() - t0
    _W["busy_s"] = sum(o[3] for o in out)   # summed single-core time of the episodes (for a per-core rate)
    return sum(o[0] for o in out), sum(o[1] for o in out), wall


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self._stop = threading.Event()
        self._t = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for k, nm in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


# --------------------------------------------------------------------------------------------- reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    cores = host_cores()
    pop = policy_population(N_EPISODES, seed=0)
    per_step = max(cores, min(4 * cores, 1024))
    for w in range(max(0, min(args.warmup, 1))):
        cpu_run(pop[:cores], cores)
    placed = evals = 0
    wall = 0.0
    for k in range(args.steps):
        sample = [pop[(k * per_step + i) % len(pop)] for i in range(per_step)]
        p, e, t = cpu_run(sample, cores)
        placed += p
        evals += e
        wall += t
    v = placed / wall
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": wall / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64+f64", "data": "bundled OpenB default trace; synthetic policy population (seed 0)",
            "config": {"workload": WORKLOAD, "note": "CPU stand-in of the reference's own path: oracle/pysim.py "
                       "(CPython + heapq, exec'd policies) on all host cores; /root/reference cannot travel to the GPU box"},
            "policy_evals_per_sec": evals / wall,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{per_step} episodes/step x {args.steps} steps of the 1024-policy population"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# --------------------------------------------------------------------------------------------- CUDA arm
def run_b200(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    from funsearch_kubernetes_simulator_b200.variants import policy_population

    # CPU baseline first (rank 0, N == 1): processes are spawned before CUDA is touched
    cpu_baseline = None
    pop = policy_population(N_EPISODES, seed=rank)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = host_cores()
        n_sample = max(cores, min(8 * cores, 128))   # about 30 s of CPU work on 16 cores
        p, e, t = cpu_run(pop[:n_sample], cores)
        cpu_baseline = {"value": p / t, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"first {n_sample} of the 1024 episodes, oracle/pysim.py on {cores} processes, {t:.1f} s",
                        "policy_evals_per_sec": e / t, "per_core_value": p / _W["busy_s"]}

    import numpy as np
    import torch
    import torch.distributed as dist
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from funsearch_kubernetes_simulator_b200 import _native as nat
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    import __graft_entry__
    if rank == 0:
        __graft_entry__.build()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.barrier()
    dev = torch.device("cuda", local)
    cluster, trace = ClusterSpec.load("gpu16"), Trace.load("default")
    ev = BatchEvaluator(cluster, [trace], device=local)

    from funsearch_kubernetes_simulator_b200 import policy_compiler
    policy_compiler._CACHE.clear()   # policy_population() already translated them once: time a cold compilation
    t0 = time.perf_counter()
    compiled = ev.compile(pop)       # translation + NVRTC, in worker processes for a population this large
    t_compile = time.perf_counter() - t0
    assert all(c is not None for c in compiled)
    t0 = time.perf_counter()
    ev.compile(pop)                  # the same sources again: translator cache + cubin cache
    t_cached = time.perf_counter() - t0
    ep_pol = list(range(N_EPISODES))
    ep_tr = [0] * N_EPISODES
    t0 = time.perf_counter()
    rb = ev.prepare_resident(compiled, ep_pol, ep_tr, teams_per_block=args.teams)
    t_link = time.perf_counter() - t0
    hbatch = ev.prepare_host_batch(compiled, ep_pol, ep_tr, teams_per_block=args.teams)

    stream = torch.cuda.current_stream(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    gathered = torch.empty(world * N_EPISODES, dtype=torch.float64, device=dev) if world > 1 else None

    def step():
        rb.launch(stream)
        if world > 1:  # the one exchange of the path: per-episode fitness to every rank
            dist.all_gather_into_tensor(gathered, rb.scores_tensor().contiguous())

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize(dev)

    launches0 = nat.lib().ks_launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kevs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local) as clocks:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        for k in range(args.steps):
            flush.fill_(k & 0xFF)  # evict L2 between timed iterations (outside the per-step events)
            evs[k][0].record(stream)
            kevs[k][0].record(stream)
            rb.launch(stream)
            kevs[k][1].record(stream)
            if world > 1:
                dist.all_gather_into_tensor(gathered, rb.scores_tensor().contiguous())
            evs[k][1].record(stream)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
    launches = nat.lib().ks_launch_count() - launches0
    step_ms = [a.elapsed_time(b) for a, b in evs]
    kern_ms = [a.elapsed_time(b) for a, b in kevs]
    res = rb.results()
    placed = int(res["placed"].sum())
    evals = int(res["scorer_calls"].sum())
    events = int(res["events"].sum())
    ok = int((res["status"] == 0).sum())
    total_ms = float(sum(step_ms))

    # end to end through the host-buffer C ABI
    for _ in range(2):
        hbatch.run(stream)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        hres = hbatch.run(stream)
        if world > 1:
            sc = torch.from_numpy(np.ascontiguousarray(hres["policy_score"])).to(dev)
            dist.all_gather_into_tensor(gathered, sc)
            torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    e2e_placed = int(hres["placed"].sum())
    assert e2e_placed == placed, "host-buffer path and resident path disagree"

    # aggregate over ranks: max time, summed work
    stats = torch.tensor([total_ms, e2e_s, float(placed), float(evals), float(events), float(sum(kern_ms))],
                         dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        total_ms, e2e_s, kern_total = float(mx[0]), float(mx[1]), float(mx[5])
        placed_all, evals_all, events_all = int(sm[2]), int(sm[3]), int(sm[4])
    else:
        kern_total = float(sum(kern_ms))
        placed_all, evals_all, events_all = placed, evals, events
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    ms_per_step = total_ms / args.steps
    value = placed_all / (ms_per_step * 1e-3)
    kern_ms_avg = kern_total / args.steps
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (measured)"
    else:
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    # per placement: the pod record; per episode: the result record and one read of the node constant rows
    alg_bytes = 32 * placed + N_EPISODES * (128 + cluster.n_nodes * 20 * 4)
    achieved = alg_bytes / (kern_ms_avg * 1e-3) / 1e9
    traffic = None
    ncu_issue = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        prof = json.load(open(tpath))
        traffic = prof.get("dram_bytes_per_launch")
        ncu_issue = {k: prof[k] for k in ("issue_slot_utilization_pct", "threads_per_warp_instruction", "icache_hit_rate_pct",
                                          "cycles_per_issued_instruction", "stall_no_instruction", "stall_wait") if k in prof}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64+f64", "data": "bundled OpenB default trace; synthetic policy population (seeded, no LLM)",
        "config": {"workload": WORKLOAD, "episodes_per_gpu": N_EPISODES, "l2": "flushed between timed steps (256 MiB write)",
                   "warps_per_team": rb.plan.warps_per_team, "teams_per_block": rb.plan.teams_per_block, "blocks": rb.plan.blocks,
                   "ctas_per_sm": rb.plan.ctas_per_sm, "smem_per_block": rb.plan.smem_per_block,
                   "parity_ok_episodes": ok},
        "policy_evals_per_sec": evals_all / (ms_per_step * 1e-3),
        "events_per_sec": events_all / (ms_per_step * 1e-3),
        "episodes_per_sec": N_EPISODES * world / (ms_per_step * 1e-3),
        "e2e": {"value": placed_all / (e2e_s / args.steps), "unit": UNIT, "h2d_bytes_per_step": hbatch.h2d_bytes,
                "d2h_bytes_per_step": hbatch.d2h_bytes, "ms_per_step": e2e_s / args.steps * 1e3,
                "path": "ks_eval_batch_host (pinned host buffers -> device workspace -> kernel -> host results)"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "kernel": "ks_replay_kernel",
                     "kernel_ms_per_launch": kern_ms_avg, "algorithmic_bytes_per_launch": alg_bytes,
                     "ncu_issue": ncu_issue,  # what actually bounds the kernel (one ncu capture, profiles/r01_f_ncu_summary.txt)
                     "note": "serial event chains bound by instruction fetch and dependent-issue latency (profiles/README.md); HBM fraction is expected << 1% (SURVEY 8d)"},
        "cpu_baseline": cpu_baseline,
        "compile": {"policies": N_EPISODES, "translate_and_nvrtc_s": t_compile, "cached_s": t_cached, "link_load_s": t_link,
                    "host_cores": host_cores(), "note": "cold population, worker processes; not part of a step"},
        "clocks": clocks.summary(),
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_RESULT_OUT = None


def emit(line):
    """The one JSON line, on the process's original stdout."""
    out = _RESULT_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    # stdout carries the JSON line only: whatever libraries print there (NCCL's version banner, worker chatter) is
    # sent to stderr by pointing fd 1 at fd 2 for the rest of the process and its children
    global _RESULT_OUT
    sys.stdout.flush()
    _RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--teams", type=int, default=0, help="warps (episodes) per CTA; 0 = planner's choice")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
