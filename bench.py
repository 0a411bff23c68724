#!/usr/bin/env python3
"""Headline benchmark: pod placements/s of FunSearch generations replayed on B200(s).

Workloads (BASELINE.json configs; --config picks one, default 1):
  1  one generation: 1,024 mutated policy variants x the bundled 8,152-pod OpenB trace x 16 nodes, one episode each,
     per GPU ("scaling": "weak"; every rank replays the SAME population unless --population rank)
  2  64 policies x a 1,000,000-pod synthetic Alibaba-shaped trace x 1,024 nodes (weak: every rank the same 64)
  3  full-population fitness: P candidates x 16 traces x 4 seeds, ONE table strong-scaled over the ranks through
     distributed.evaluate_episodes_sharded (LPT on predicted cost) with the NCCL fitness all-gather
  4  episode-count sweep 2^10 .. 2^17 episodes of the config-1 shape (one JSON line per size)

One "step" = one launch of the replay kernel over the whole batch (+ the NCCL all-gather of per-episode fitness
when N > 1).

  python bench.py [--config C] [--gpus N] [--steps K] [--warmup W]      # this framework (CUDA)
  python bench.py --impl reference [--config C] [--steps K] [--warmup W]  # CPU stand-in of the reference
  torchrun ... bench.py --gpus N ...                                    # one rank per GPU

Prints ONE JSON line (rank 0; config 4: one per size).  `value` = placements/s with inputs resident in HBM (CUDA
events on the launch stream); `e2e` = the same metric through the host-buffer C-ABI call ks_eval_batch_host with the
workload copied host->device and results device->host inside the timed region; `e2e_fresh` = wall time of
BatchEvaluator.evaluate_population on sources this process has never seen (translator, NVRTC, link, module load,
launch, read-back all inside); `roofline` relates the algorithmic bytes (32 B per placement) to the
measured HBM peak; `cpu_baseline` is the oracle port (oracle/pysim.py: same algorithm, CPython + heapq like the
reference) on the host cores, and its episodes are compared with the GPU's (`parity_checked_episodes`).
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
WORKLOADS = {
    1: "1024 mutated policy variants x default OpenB trace (8152 pods) x 16 nodes, 1 episode each (configs[1])",
    2: "64 policies x synthetic Alibaba-shaped trace (1,000,000 pods, load factor 200) x 1024 nodes (configs[2])",
    3: "population x 16 bundled traces x 4 bootstrap seeds x 16 nodes, one table strong-scaled over the GPUs (configs[3])",
    4: "episode-count sweep 2^10..2^17 episodes of the configs[1] shape (configs[4])",
}


# config 2: the six shipped policies the C oracle has hand translations of ride along with the 58 variants, so the measured
# batch itself is checked against the CPU (name in tests/golden/policies.json -> built-in of oracle/ksim_oracle.c)
C_ORACLE_POLICIES = {"ts_first_fit": "first_fit", "ts_best_fit": "best_fit", "ts_funsearch_4901": "funsearch_4901",
                     "ts_funsearch_4816": "funsearch_4816", "ts_funsearch_4800": "funsearch_4800", "fs_worst_fit": "worst_fit"}
CONFIG2_EVENT_BUDGET = 4000000   # per episode: a spreading policy never catches up with this arrival rate (status 3)


def workload_config(args):
    """The `config` object: identical in both arms (b200 / reference) for the same command line."""
    c = {"workload": WORKLOADS[args.config], "config_id": args.config, "population": f"variants.policy_population(seed={args.seed})",
         "population_per_rank": "same" if args.population == "same" else "seed + rank"}
    if args.config == 3:
        c["candidates"] = args.candidates
    if args.config == 2:
        c["policies_per_gpu"] = args.policies_per_gpu
        c["workload"] = c["workload"].replace("64 policies", f"{args.policies_per_gpu} policies per GPU")
    return c


# --------------------------------------------------------------------------------------------- CPU side
_W = {}


def _cpu_init(cluster_name, n_nodes, trace_name, head):
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_specs
    cluster = ClusterSpec.load(cluster_name)
    if n_nodes:
        cluster = cluster.head(n_nodes)
    trace = Trace.load(trace_name)
    if head:
        trace = trace.head(head)
    _W["specs"] = oracle_specs(cluster, trace)


def _cpu_episode(src):
    from oracle import pysim
    cs, ts = _W["specs"]
    t0 = time.perf_counter()
    try:
        r = pysim.run_episode(cs, ts, pysim.scorer_from_source(src))
    except Exception:  # noqa: BLE001  (a source that does not even define the function: the worker scores it 0)
        return 0, 0, 0, time.perf_counter() - t0, (1, 0, 0, 0, 0.0)
    return r.placed, r.scorer_calls, r.events, time.perf_counter() - t0, (r.status, r.events, r.attempts, r.placed, r.score)


def cpu_run(sources, workers, spec=("gpu16", 0, "default", 0)):
    """Replays `sources` on one workload with the oracle port on `workers` processes.
    Returns (placements, scorer_calls, wall seconds, per-episode (status, events, attempts, placed, score))."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    with ctx.Pool(workers, initializer=_cpu_init, initargs=spec) as pool:
        pool.map(_cpu_episode, sources[:workers])  # warm the workers (imports, trace load)
        t0 = time.perf_counter()
        out = pool.map(_cpu_episode, sources, chunksize=1)
        wall = time.perf_counter() - t0
    _W["busy_s"] = sum(o[3] for o in out)   # summed single-core time of the episodes (for a per-core rate)
    return sum(o[0] for o in out), sum(o[1] for o in out), wall, [o[4] for o in out]


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


# --------------------------------------------------------------------------------------------- workloads
def build_workload(args, rank, world):
    """(cluster, traces, sources, ep_policy, ep_trace, cpu_spec) for this rank.  Weak-scaled configs give every rank
    its own full table; config 3 gives every rank the WHOLE table (the sharded API cuts it)."""
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace, bootstrap_trace, synthetic_trace
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    seed = args.seed + (rank if args.population == "rank" else 0)
    if args.config in (1, 4):
        cluster, traces = ClusterSpec.load("gpu16"), [Trace.load("default")]
        n = args.episodes or 1024
        pop = policy_population(min(n, 1024), seed=seed)
        ep_pol = [i % len(pop) for i in range(n)]
        ep_tr = [0] * n
        return cluster, traces, pop, ep_pol, ep_tr, ("gpu16", 0, "default", 0)
    if args.config == 2:
        cluster = ClusterSpec.load("all1523").head(1024)
        traces = [synthetic_trace(args.pods or 1000000, seed=3, load_factor=200.0)]
        shipped = json.load(open(os.path.join(ROOT, "tests", "golden", "policies.json")))
        npol = args.policies_per_gpu
        pop = policy_population(npol - len(C_ORACLE_POLICIES), seed=seed) + [shipped[k] for k in C_ORACLE_POLICIES]
        return cluster, traces, pop, list(range(npol)), [0] * npol, None
    if args.config == 3:
        cluster = ClusterSpec.load("gpu16")
        names = ["default", "cpu037", "cpu050", "cpu072", "cpu100", "cpu200", "cpu235", "cpu250", "cpu300", "gpushare20",
                 "gpushare40", "gpushare60", "gpushare80", "gpushare100", "gpuspec10", "gpuspec25"]
        traces = [bootstrap_trace(Trace.load(nm), s) for nm in names for s in range(4)]
        pop = policy_population(args.candidates, seed=args.seed)
        ep_pol = [p for p in range(len(pop)) for _t in range(len(traces))]
        ep_tr = [t for _p in range(len(pop)) for t in range(len(traces))]
        return cluster, traces, pop, ep_pol, ep_tr, None
    raise ValueError(args.config)


def config2_cpu_column(cluster, trace, pop):
    """CPU column of config 2: the C oracle (oracle/ksim_oracle.c, -O2, the same algorithm with hand-translated policies)
    replays the six shipped policies on the full trace, one thread each; plus the Python port on a prefix of the trace,
    extrapolated (the Python path needs minutes per episode at this size)."""
    import threading
    from oracle import coracle, pysim
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_specs
    names = list(C_ORACLE_POLICIES)
    out = {}

    def work(nm):
        t0 = time.perf_counter()
        res, _t, _n, _m = coracle.run(cluster, trace, C_ORACLE_POLICIES[nm], max_events=CONFIG2_EVENT_BUDGET)
        out[nm] = (res.status, res.events, res.attempts, res.placed, res.score, time.perf_counter() - t0)
    coracle.lib()
    ths = [threading.Thread(target=work, args=(nm,)) for nm in names]
    t0 = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    wall = time.perf_counter() - t0
    placed = sum(v[3] for v in out.values())
    busy = sum(v[5] for v in out.values())
    # Python port: first 2,000 pods of the trace under first-fit, one core
    n_prefix = 2000
    cs, ts = oracle_specs(cluster, trace.head(n_prefix))
    shipped = json.load(open(os.path.join(ROOT, "tests", "golden", "policies.json")))
    t0 = time.perf_counter()
    r = pysim.run_episode(cs, ts, pysim.scorer_from_source(shipped["ts_first_fit"]))
    t_py = time.perf_counter() - t0
    cores = host_cores()
    col = {"value": placed / busy * cores, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"C oracle (gcc -O2), the 6 shipped policies x the full trace (event budget {CONFIG2_EVENT_BUDGET}), "
                     f"one thread each: {placed} placements in {busy:.1f} core-seconds ({wall:.1f} s wall); value = per-core rate x {cores} cores",
           "per_core_value": placed / busy,
           "python_port_extrapolated": {"per_core_value": r.placed / t_py, "value": r.placed / t_py * cores,
                                        "sample": f"oracle/pysim.py, first-fit, first {n_prefix} pods of the trace x 1,024 nodes on one core "
                                                  f"({t_py:.1f} s), extrapolated to {cores} cores; labelled extrapolation (SURVEY 8d)"}}
    return col, {nm: v[:5] for nm, v in out.items()}


# --------------------------------------------------------------------------------------------- reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    cores = host_cores()
    if args.config != 1:
        emit({"impl": "reference", "metric": METRIC, "unit": UNIT, "config": workload_config(args),
              "unavailable": "the CPU arm is timed on configs[1]; configs 2-4 carry their own bounded CPU column in cpu_baseline"})
        return
    pop = policy_population(1024, seed=args.seed)
    per_step = max(cores, min(4 * cores, 1024))
    for w in range(max(0, min(args.warmup, 1))):
        cpu_run(pop[:cores], cores)
    placed = evals = 0
    wall = 0.0
    for k in range(args.steps):
        sample = [pop[(k * per_step + i) % len(pop)] for i in range(per_step)]
        p, e, t, _ = cpu_run(sample, cores)
        placed += p
        evals += e
        wall += t
    v = placed / wall
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": wall / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64+f64", "data": "bundled OpenB default trace; synthetic policy population (seeded, no LLM)",
            "config": workload_config(args),
            "note": "CPU stand-in of the reference's own path: oracle/pysim.py (CPython + heapq, exec'd policies) on all host "
                    "cores; /root/reference cannot travel to the GPU box (BASELINE.md: the port runs 1.28x faster than the reference)",
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
    t_gen = time.perf_counter()
    cluster, traces, pop, ep_pol, ep_tr, cpu_spec = build_workload(args, rank, world)
    t_gen = time.perf_counter() - t_gen
    n_ep = len(ep_pol)

    # CPU baseline first (rank 0, N == 1): processes are spawned before CUDA is touched.  Its episodes are kept:
    # they are compared with the GPU's results for the same (policy, trace) pairs below.
    cpu_baseline = None
    cpu_rows = []
    n_sample = 0
    if rank == 0 and world == 1 and not args.no_cpu_baseline and cpu_spec is not None:
        cores = host_cores()
        n_sample = min(n_ep, max(cores, min(8 * cores, 128)))   # about 30 s of CPU work on 16 cores
        p, e, t, cpu_rows = cpu_run([pop[ep_pol[i]] for i in range(n_sample)], cores, cpu_spec)
        cpu_baseline = {"value": p / t, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"first {n_sample} of the {n_ep} episodes, oracle/pysim.py on {cores} processes, {t:.1f} s",
                        "policy_evals_per_sec": e / t, "per_core_value": p / _W["busy_s"]}

    c2_rows = {}
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.config == 2:
        cpu_baseline, c2_rows = config2_cpu_column(cluster, traces[0], pop)

    import numpy as np
    import torch
    import torch.distributed as dist
    from funsearch_kubernetes_simulator_b200 import _native as nat
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    from funsearch_kubernetes_simulator_b200 import compile_pool
    import __graft_entry__
    if rank == 0:
        __graft_entry__.build()
    torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if world > 1:
        dist.barrier()
    dev = torch.device("cuda", local)
    t0 = time.perf_counter()
    ev = BatchEvaluator(cluster, traces, device=local)
    pool_info = compile_pool.wait_ready() if compile_pool.n_workers() else {"workers": 0, "warmup_s_max": 0.0}
    t_pool = time.perf_counter() - t0

    from funsearch_kubernetes_simulator_b200 import policy_compiler
    policy_compiler._CACHE.clear()   # the generator does not translate, but be explicit: time a cold compilation
    t0 = time.perf_counter()
    compiled = ev.compile(pop)       # translation + NVRTC of the new structures, in worker processes
    t_compile = time.perf_counter() - t0
    cstats = dict(ev.last_compile_stats)
    untranslated = [i for i, c in enumerate(compiled) if c is None]
    assert not untranslated, f"{len(untranslated)} sources of the population do not translate, e.g. #{untranslated[0]}: " \
                             f"{ev.compile_error(pop[untranslated[0]])}"
    t0 = time.perf_counter()
    ev.compile(pop)                  # the same sources again: translator cache + registered structures
    t_cached = time.perf_counter() - t0
    t0 = time.perf_counter()
    single = True if args.single_module else None
    if not args.first_evaluation_modules:
        # Steady state: the compile workers hand a generation over as one module per worker; between generations the
        # evaluator relinks it into ONE module in the background (evaluator._consolidate_async).  The timed steps replay
        # the generation as it is replayed from then on; e2e_fresh below pays for the per-worker modules instead.
        ev.bind(compiled)
        ev.wait_background()
    budget = CONFIG2_EVENT_BUDGET if args.config == 2 else 0
    rb = ev.prepare_resident(compiled, ep_pol, ep_tr, teams_per_block=args.teams, single_module=single, max_events=budget)
    t_link = time.perf_counter() - t0
    hbatch = ev.prepare_host_batch(compiled, ep_pol, ep_tr, teams_per_block=args.teams, single_module=single, max_events=budget)

    stream = torch.cuda.current_stream(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    gathered = torch.empty(world * n_ep, dtype=torch.float64, device=dev) if world > 1 else None

    def step():
        rb.launch(stream)
        if world > 1:  # the one exchange of the path: per-episode fitness to every rank
            dist.all_gather_into_tensor(gathered, rb.scores_tensor().contiguous())

    # (config 2: one step replays 64 x up to 4 M strictly serial events, about 20 s: the caller's warm-up count is taken
    #  as is there; everywhere else at least three warm-up steps)
    n_warm = max(args.warmup, 1) if args.config == 2 else max(args.warmup, 3)
    for _ in range(n_warm):
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
    cyc = res["device_cycles"].astype(np.float64)

    # parity of the measured workload: the CPU leg's episodes against the GPU's
    parity_checked = parity_bad = parity_declined = 0
    first_bad = None
    for i, row in enumerate(cpu_rows):
        g = res[i]
        got = (int(g["status"]), int(g["events"]), int(g["attempts"]), int(g["placed"]))
        if got[0] == 4:   # the GPU left the translated subset (complex number, int beyond int64): reported, not compared
            parity_declined += 1
            continue
        same = got == tuple(row[:4]) and abs(float(g["policy_score"]) - row[4]) <= 1e-9 * max(abs(row[4]), 1e-300)
        if row[0] != 0:   # a failed episode: the status and a zero score are what must agree
            same = got[0] == row[0] and float(g["policy_score"]) == 0.0
        parity_checked += 1
        if not same:
            parity_bad += 1
            if first_bad is None:
                first_bad = {"episode": i, "gpu": got + (float(g["policy_score"]),), "cpu": list(row)}

    for nm, row in c2_rows.items():   # config 2: the shipped policies of the batch against the C oracle
        i = len(pop) - len(C_ORACLE_POLICIES) + list(C_ORACLE_POLICIES).index(nm)
        g = res[i]
        got = (int(g["status"]), int(g["events"]), int(g["attempts"]), int(g["placed"]))
        same = got == tuple(row[:4]) and abs(float(g["policy_score"]) - row[4]) <= 1e-9 * max(abs(row[4]), 1e-300)
        parity_checked += 1
        if not same:
            parity_bad += 1
            if first_bad is None:
                first_bad = {"episode": i, "policy": nm, "gpu": got + (float(g["policy_score"]),), "cpu": list(row)}

    # end to end through the host-buffer C ABI
    for _ in range(1 if args.config == 2 else 2):
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

    # a fresh generation end to end through the reference-shaped call: sources never seen by this process
    fresh = None
    if args.fresh_steps > 0 and args.config == 1:
        from funsearch_kubernetes_simulator_b200.variants import policy_population
        gens = [policy_population(1024, seed=100003 + 7919 * (k + 1) + rank) for k in range(args.fresh_steps + 1)]
        walls, stats, placed_f, structs, not_eval = [], [], [], [], []
        for k, g in enumerate(gens):
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            ev.evaluate_population(g)           # translate + NVRTC + link + load + launch + read-back
            dt = time.perf_counter() - t0
            if k == 0:
                continue                        # untimed warm-up generation (first use of the linker etc.)
            walls.append(dt)
            stats.append(dict(ev.last_compile_stats, kernel_ms=ev.last_kernel_ms))
            structs.append(ev.last_report.distinct_structures)
            not_eval.append(len(ev.last_report.not_evaluated))
            placed_f.append(8152 * sum(1 for s in ev.last_report.status if s == 0))
        fresh = {"value": sum(placed_f) / sum(walls), "unit": UNIT, "ms_per_generation": [round(w * 1e3, 1) for w in walls],
                 "generations": len(walls), "policies_per_generation": 1024, "distinct_structures": structs,
                 "not_evaluated": not_eval, "stages": stats,
                 "path": "BatchEvaluator.evaluate_population(sources): translate + NVRTC of new structures (worker pool) + "
                         "nvJitLink + module load + H2D + kernel + D2H; one untimed warm-up generation before"}

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
        return

    ms_per_step = total_ms / args.steps
    value = placed_all / (ms_per_step * 1e-3)
    kern_ms_avg = kern_total / args.steps
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (measured)"
    else:
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
    # per placement: the pod record + placement record; per episode: the result record and one read of the node constant rows
    alg_bytes = 32 * placed + n_ep * (128 + cluster.n_nodes * 20 * 4)
    achieved = alg_bytes / (kern_ms_avg * 1e-3) / 1e9
    traffic = None
    ncu_issue = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath) and args.config == 1:
        prof = json.load(open(tpath))
        traffic = prof.get("dram_bytes_per_launch")
        ncu_issue = {k: prof[k] for k in ("issue_slot_utilization_pct", "threads_per_warp_instruction", "icache_hit_rate_pct",
                                          "cycles_per_issued_instruction", "stall_no_instruction", "stall_wait") if k in prof}
    cfg = workload_config(args)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": n_warm,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64+f64", "data": "bundled OpenB default trace; synthetic policy population (seeded, no LLM)",
        "config": cfg,
        "launch": {"episodes_per_gpu": n_ep, "l2": "flushed between timed steps (256 MiB write)",
                   "warps_per_team": rb.plan.warps_per_team, "teams_per_block": rb.plan.teams_per_block, "blocks": rb.plan.blocks,
                   "ctas_per_sm": rb.plan.ctas_per_sm, "smem_per_block": rb.plan.smem_per_block,
                   "distinct_structures": rb.n_structures, "distinct_sources": len(set(pop)),
                   "status_ok_episodes": ok,
                   "slowest_episode_mcycles": float(cyc.max() / 1e6), "mean_episode_mcycles": float(cyc.mean() / 1e6)},
        "parity_checked_episodes": parity_checked, "parity_mismatches": parity_bad, "parity_first_mismatch": first_bad,
        "parity_gpu_declined_episodes": parity_declined,   # status 4 'unsupported' on the GPU: not evaluated, not compared
        "policy_evals_per_sec": evals_all / (ms_per_step * 1e-3),
        "events_per_sec": events_all / (ms_per_step * 1e-3),
        "episodes_per_sec": n_ep * world / (ms_per_step * 1e-3),
        "e2e": {"value": placed_all / (e2e_s / args.steps), "unit": UNIT, "h2d_bytes_per_step": hbatch.h2d_bytes,
                "d2h_bytes_per_step": hbatch.d2h_bytes, "ms_per_step": e2e_s / args.steps * 1e3,
                "path": "ks_eval_batch_host (pinned host buffers -> device workspace -> kernel -> host results)"},
        "e2e_fresh": fresh,
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic,
                     "traffic_source": "replayed from profiles/traffic.json (one ncu --set full capture of this command by the builder), not measured in this run" if traffic else None,
                     "peak_source": peak_src, "kernel": "ks_replay_kernel",
                     "kernel_ms_per_launch": kern_ms_avg, "algorithmic_bytes_per_launch": alg_bytes,
                     "ncu_issue": ncu_issue,
                     "note": "serial event chains bound by instruction fetch and dependent-issue latency (profiles/README.md); HBM fraction is expected << 1% (SURVEY 8d)"},
        "cpu_baseline": cpu_baseline,
        "compile": {"policies": len(pop), "distinct_structures": cstats.get("new_structures"), "total_s": t_compile,
                    "translate_s": cstats.get("translate_s"), "nvrtc_link_s": cstats.get("nvrtc_link_s"),
                    "register_load_s": cstats.get("register_load_s"), "modules_loaded": cstats.get("modules_loaded"),
                    "worker_translate_s_max": cstats.get("worker_translate_s_max"),
                    "worker_nvrtc_link_s_max": cstats.get("worker_nvrtc_link_s_max"),
                    "cached_s": t_cached, "bind_s": t_link, "modules_in_batch": rb.n_modules, "pool_workers": pool_info["workers"],
                    "pool_warmup_s_max": pool_info["warmup_s_max"], "evaluator_and_pool_ready_s": t_pool,
                    "population_generation_s": t_gen, "host_cores": host_cores(),
                    "note": "cold population, worker processes; not part of a step (e2e_fresh is the number that includes it)"},
        "clocks": clocks.summary(),
    }
    emit(line)


# --------------------------------------------------------------------------------------------- config 3: strong scaling
def run_config3(args):
    """One population x 16 traces x 4 bootstrap seeds, ONE episode table cut over the ranks by policy (LPT on each
    policy's predicted cost), through the product's sharded API; the single exchange is the all-gather of per-policy
    fitness.  `value`: launch + device-side per-policy mean + NCCL all-gather, inputs resident.  `e2e`:
    distributed.evaluate_population_sharded(sources, BatchEvaluator.evaluate_population, costs) wall clock."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import numpy as np
    import torch
    import torch.distributed as dist
    from funsearch_kubernetes_simulator_b200 import _native as nat
    from funsearch_kubernetes_simulator_b200 import compile_pool
    from funsearch_kubernetes_simulator_b200.distributed import evaluate_population_sharded, shard_plan
    from funsearch_kubernetes_simulator_b200.evaluator import BatchEvaluator
    import __graft_entry__
    cluster, traces, pop, _ep, _et, _ = build_workload(args, rank, world)
    if rank == 0:
        __graft_entry__.build()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.barrier()
    dev = torch.device("cuda", local)
    ev = BatchEvaluator(cluster, traces, device=local)
    if compile_pool.n_workers():
        compile_pool.wait_ready()
    n_tr = len(traces)
    t0 = time.perf_counter()
    compiled_all = ev.compile(pop)
    t_compile = time.perf_counter() - t0
    keep = [i for i, c in enumerate(compiled_all) if c is not None]
    pop = [pop[i] for i in keep]
    compiled_all = [compiled_all[i] for i in keep]
    # predicted cost of a policy = the events it generated on the first trace (in an evolution run: last generation's)
    cal = ev.prepare_resident(compiled_all, list(range(len(pop))), [0] * len(pop))
    cal.launch()
    costs = [float(x) for x in cal.results()["events"]]
    plan = shard_plan(len(pop), world, costs)
    mine = plan[rank]
    cap = max(len(sh) for sh in plan)
    my_pol = [compiled_all[i] for i in mine]
    ep_pol = [p for p in range(len(mine)) for _t in range(n_tr)]
    ep_tr = [t for _p in range(len(mine)) for t in range(n_tr)]
    rb = ev.prepare_resident(my_pol, ep_pol, ep_tr)
    stream = torch.cuda.current_stream(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    block = torch.zeros(cap, dtype=torch.float64, device=dev)
    gathered = torch.empty(world * cap, dtype=torch.float64, device=dev)

    def step():
        rb.launch(stream)
        sc = rb.scores_tensor().view(len(mine), n_tr)
        ok = (rb.status_tensor().view(len(mine), n_tr) == 0).all(dim=1)
        block[:len(mine)] = torch.where(ok, sc.mean(dim=1), torch.zeros_like(sc[:, 0]))
        if world > 1:
            dist.all_gather_into_tensor(gathered, block)
    for _ in range(max(1, min(args.warmup, 3))):
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
            flush.fill_(k & 0xFF)
            evs[k][0].record(stream)
            kevs[k][0].record(stream)
            rb.launch(stream)
            kevs[k][1].record(stream)
            sc = rb.scores_tensor().view(len(mine), n_tr)
            ok = (rb.status_tensor().view(len(mine), n_tr) == 0).all(dim=1)
            block[:len(mine)] = torch.where(ok, sc.mean(dim=1), torch.zeros_like(sc[:, 0]))
            if world > 1:
                dist.all_gather_into_tensor(gathered, block)
            evs[k][1].record(stream)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
    launches = nat.lib().ks_launch_count() - launches0
    total_ms = float(sum(a.elapsed_time(b) for a, b in evs))
    kern_ms = float(sum(a.elapsed_time(b) for a, b in kevs))
    res = rb.results()
    placed, evals, events = int(res["placed"].sum()), int(res["scorer_calls"].sum()), int(res["events"].sum())
    device_fitness = block[:len(mine)].cpu().numpy().copy()
    # end to end through the sharded product API (cached compilation; host tables in, fitness list out)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    fit = evaluate_population_sharded(pop, ev.evaluate_population, costs, device=dev)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    e2e_s = time.perf_counter() - t0
    # (the device-side mean and the host-side mean add the 64 traces in different orders: compared to 1e-12, not bit for bit)
    mism = int(sum(1 for k, i in enumerate(mine) if abs(fit[i] - float(device_fitness[k])) > 1e-12))
    stats = torch.tensor([total_ms, e2e_s, float(placed), float(evals), float(events), kern_ms, float(mism)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        mn = stats.clone()
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        total_ms, e2e_s, kern_max, kern_min = float(mx[0]), float(mx[1]), float(mx[5]), float(mn[5])
        placed_all, evals_all, events_all, mism = int(sm[2]), int(sm[3]), int(sm[4]), int(sm[6])
    else:
        kern_max = kern_min = kern_ms
        placed_all, evals_all, events_all = placed, evals, events
    if rank == 0:
        ms_per_step = total_ms / args.steps
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak = json.load(open(peaks_path))["hbm_gbs"] if os.path.exists(peaks_path) else 6650.0
        alg_bytes = 32 * placed + len(ep_pol) * (128 + cluster.n_nodes * 80)
        achieved = alg_bytes / (kern_max / args.steps * 1e-3) / 1e9
        emit({"metric": METRIC, "value": placed_all / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
              "warmup": max(1, min(args.warmup, 3)), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
              "vs_baseline": None, "dtype": "int64+f64",
              "data": "16 bundled OpenB traces x 4 bootstrap resamples; synthetic policy population (seeded, no LLM)",
              "config": workload_config(args),
              "launch": {"episodes_total": len(pop) * n_tr, "episodes_this_rank": len(ep_pol), "policies": len(pop), "traces": n_tr,
                         "sharding": "by policy, LPT on events of a calibration episode (distributed.shard_plan)",
                         "largest_shard_policies": cap, "ctas_per_sm": rb.plan.ctas_per_sm, "teams_per_block": rb.plan.teams_per_block,
                         "blocks": rb.plan.blocks, "modules_in_batch": rb.n_modules, "l2": "flushed between timed steps (256 MiB write)",
                         "kernel_ms_slowest_rank": kern_max / args.steps, "kernel_ms_fastest_rank": kern_min / args.steps,
                         "limiter": "slowest shard's kernel (step - slowest kernel = reduction + all-gather)",
                         "step_minus_slowest_kernel_ms": ms_per_step - kern_max / args.steps},
              "parity_checked_episodes": 0, "parity_mismatches": 0,
              "sharded_api_vs_device_fitness_mismatches": mism,
              "policy_evals_per_sec": evals_all / (ms_per_step * 1e-3), "events_per_sec": events_all / (ms_per_step * 1e-3),
              "episodes_per_sec": len(pop) * n_tr / (ms_per_step * 1e-3),
              "e2e": {"value": placed_all / e2e_s, "unit": UNIT, "ms_per_step": e2e_s * 1e3,
                      "h2d_bytes_per_step": int(len(ep_pol) * 12), "d2h_bytes_per_step": int(len(ep_pol) * 128),
                      "path": "distributed.evaluate_population_sharded(sources, BatchEvaluator.evaluate_population, costs): shard, "
                              "episode tables H2D, kernel, results D2H, NCCL all-gather of fitness (one call, compilation cached)"},
              "gpu_launches": int(launches),
              "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                           "kernel": "ks_replay_kernel_dense" if rb.plan.ctas_per_sm == 2 else "ks_replay_kernel",
                           "note": "this rank's launch; serial event chains, HBM fraction expected << 1% (SURVEY 8d)"},
              "cpu_baseline": None, "compile": {"policies": len(pop), "total_s": t_compile},
              "clocks": clocks.summary()})


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
    ap.add_argument("--config", type=int, default=1, choices=[1, 2, 3, 4])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--population", default="same", choices=["same", "rank"],
                    help="weak scaling: every rank replays the same population (default) or its own (seed + rank)")
    ap.add_argument("--episodes", type=int, default=0, help="configs 1/4: episodes per GPU (default 1024)")
    ap.add_argument("--pods", type=int, default=0, help="config 2: pods of the synthetic trace (default 1,000,000)")
    ap.add_argument("--candidates", type=int, default=1024, help="config 3: policies in the population")
    ap.add_argument("--policies-per-gpu", type=int, default=64,
                    help="config 2: policies (= episodes) per GPU; 64 is configs[2], 128 is the north-star shape (1,024 policies on 8 GPUs)")
    ap.add_argument("--fresh-steps", type=int, default=2, help="timed fresh generations for e2e_fresh (0 = skip)")
    ap.add_argument("--teams", type=int, default=0, help="warps (episodes) per CTA; 0 = planner's choice")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--single-module", action="store_true", help="relink the batch's structures into one module (one launch)")
    ap.add_argument("--first-evaluation-modules", action="store_true",
                    help="time the steps on the per-worker modules of the first evaluation instead of the consolidated module")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.config == 3:
        run_config3(args)
    elif args.config == 4:
        first = True
        for lg in range(10, 18):
            args.episodes = 1 << lg
            args.fresh_steps = 0
            args.no_cpu_baseline = args.no_cpu_baseline or not first
            run_b200(args)
            first = False
    else:
        run_b200(args)
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        pass


if __name__ == "__main__":
    main()
