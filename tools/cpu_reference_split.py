"""CPU baseline split of SURVEY.md section 8d, measured where the real reference is importable (the build container,
/root/reference; it cannot travel to the GPU box): the reference's own timed region (run_schedule only,
tests/test_scheduler.py:266-269) and the full cost of one pool worker (parse_workload + deepcopy + run_schedule,
funsearch/funsearch_integration.py:36-59), on one core and on min(nproc, 8) workers (configs/llm_config.json:23), next to
the oracle port (oracle/pysim.py) on the same policies -- the ratio that relates bench.py's CPU arm to the reference.
Writes profiles/cpu_reference_split_r02.json."""
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
N_POLICIES = 16


def _ref_setup():
    os.environ["PYTHONDONTWRITEBYTECODE"] = "1"
    os.chdir(REF)
    sys.path.insert(0, REF)


def ref_episode(args):
    """One reference episode.  Returns (run_schedule seconds, full worker seconds, placed, score)."""
    code, full = args
    _ref_setup()
    import copy
    from benchmarks.parser import TraceParser
    from funsearch.funsearch_integration import FunSearchScheduler
    from funsearch.safe_execution import SafeExecutor
    from simulator.evaluator import SchedulingEvaluator
    from simulator.event_simulator import DiscreteEventSimulator
    from simulator.main import KubernetesSimulator
    t_full = time.perf_counter()
    if full or "w" not in _CACHE:
        cluster0, pods0 = TraceParser().parse_workload()
        _CACHE["w"] = (cluster0, pods0)
    cluster0, pods0 = _CACHE["w"]
    cluster, pods = copy.deepcopy(cluster0), copy.deepcopy(pods0)
    sched = FunSearchScheduler(code, SafeExecutor(3))
    sim = KubernetesSimulator(cluster, pods, DiscreteEventSimulator(pods), sched, evaluator=SchedulingEvaluator(cluster, enabled=True))
    t0 = time.perf_counter()
    try:
        sim.run_schedule()
        score = sim.evaluator.get_policy_score(pods)
    except Exception:  # noqa: BLE001
        score = 0
    t1 = time.perf_counter()
    return t1 - t0, t1 - t_full, sum(1 for p in pods if p.assigned_node != ""), score


_CACHE = {}


def port_episode(code):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_specs
    from funsearch_kubernetes_simulator_b200 import ClusterSpec, Trace
    from oracle import pysim
    if "s" not in _CACHE:
        _CACHE["s"] = oracle_specs(ClusterSpec.load("gpu16"), Trace.load("default"))
    cs, ts = _CACHE["s"]
    t0 = time.perf_counter()
    r = pysim.run_episode(cs, ts, pysim.scorer_from_source(code))
    return time.perf_counter() - t0, r.placed, r.score


def main():
    sys.path.insert(0, ROOT)
    from funsearch_kubernetes_simulator_b200.variants import policy_population
    pop = policy_population(1024, seed=0)[:N_POLICIES]
    workers = min(os.cpu_count() or 1, 8)
    ctx = mp.get_context("spawn")
    out = {"policies": N_POLICIES, "population": "variants.policy_population(1024, seed=0)[:16]", "workers": workers,
           "host": os.uname().machine, "python": sys.version.split()[0]}
    with ctx.Pool(1) as pool:
        pool.map(ref_episode, [(pop[0], False)])
        r = pool.map(ref_episode, [(c, False) for c in pop])
        rf = pool.map(ref_episode, [(c, True) for c in pop])
    with ctx.Pool(1) as pool:
        pool.map(port_episode, pop[:1])
        p = pool.map(port_episode, pop)
    assert [x[2] for x in r] == [x[1] for x in p] and all(abs(a[3] - b[2]) < 1e-15 for a, b in zip(r, p)), "port != reference"
    placed = sum(x[2] for x in r)
    out["single_core"] = {
        "reference_run_schedule_s": sum(x[0] for x in r), "reference_full_worker_s": sum(x[1] for x in rf),
        "port_run_schedule_s": sum(x[0] for x in p), "placements": placed,
        "reference_run_schedule_placements_per_s": placed / sum(x[0] for x in r),
        "reference_full_worker_placements_per_s": placed / sum(x[1] for x in rf),
        "port_placements_per_s": placed / sum(x[0] for x in p),
        "port_speed_over_reference": sum(x[0] for x in r) / sum(x[0] for x in p)}
    with ctx.Pool(workers) as pool:
        pool.map(ref_episode, [(pop[0], False)] * workers)
        t0 = time.perf_counter()
        pool.map(ref_episode, [(c, False) for c in pop], chunksize=1)
        t_rs = time.perf_counter() - t0
        t0 = time.perf_counter()
        pool.map(ref_episode, [(c, True) for c in pop], chunksize=1)
        t_full = time.perf_counter() - t0
    out[f"workers_{workers}"] = {"reference_run_schedule_wall_s": t_rs, "reference_full_worker_wall_s": t_full,
                                 "reference_run_schedule_placements_per_s": placed / t_rs,
                                 "reference_full_worker_placements_per_s": placed / t_full,
                                 "note": "wall clock of the pool over the 16 episodes; the first includes each worker's deepcopy but "
                                         "not parse_workload, the second re-parses the CSVs per episode like the reference's worker"}
    path = os.path.join(ROOT, "profiles", "cpu_reference_split_r02.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
