#!/usr/bin/env python3
"""Regenerates the committed fixtures from the REAL reference (build container only).

Run:  python tests/golden/make_golden.py            (needs /root/reference; no GPU)

It (1) converts the reference's bundled OpenB CSV traces into the compact .npz workload
files under funsearch_kubernetes_simulator_b200/data/, (2) collects the shipped policy
sources into tests/golden/policies.json, (3) runs the reference's own
KubernetesSimulator and oracle/pysim.py side by side on every usable trace x policy and
ASSERTS they agree on placement log, counters and metrics, and (4) writes
tests/golden/golden_runs.json.  Nothing in tests/, bench.py or smoke() reads
/root/reference at run time; they read what this script wrote.
"""
import copy
import hashlib
import inspect
import json
import os
import platform
import random
import re
import sys
import time

import numpy as np

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
REF = "/root/reference"
sys.dont_write_bytecode = True
sys.path.insert(0, REPO)
sys.path.insert(0, REF)
os.chdir(REF)  # the reference parser uses a relative traces dir (benchmarks/parser.py:12)

from benchmarks.parser import TraceParser  # noqa: E402
from simulator.event_simulator import DiscreteEventSimulator  # noqa: E402
from simulator.evaluator import SchedulingEvaluator  # noqa: E402
from simulator.main import KubernetesSimulator  # noqa: E402
from simulator.entities import Cluster, Node, GPU, Pod  # noqa: E402
from funsearch.funsearch_integration import FunSearchScheduler, SimpleFunSearch  # noqa: E402
from funsearch.safe_execution import SafeExecutor, PolicyTemplate  # noqa: E402
import importlib.util  # noqa: E402


def _load_ref_module(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


ref_ts = _load_ref_module("ref_test_scheduler", os.path.join(REF, "tests", "test_scheduler.py"))

from oracle import pysim  # noqa: E402

DATA = os.path.join(REPO, "funsearch_kubernetes_simulator_b200", "data")
GOLD = os.path.join(REPO, "tests", "golden")


# ----------------------------------------------------------------------------- workloads
def cluster_to_spec(cluster) -> pysim.ClusterSpec:
    ids, cpu, mem, gpus = [], [], [], []
    for nid, n in cluster.nodes_dict.items():
        ids.append(nid)
        cpu.append(n.cpu_milli_total)
        mem.append(n.memory_mib_total)
        gpus.append([(g.memory_mib_total, g.gpu_milli_total) for g in n.gpus])
    return pysim.ClusterSpec(ids, cpu, mem, gpus)


def pods_to_spec(pods) -> pysim.TraceSpec:
    return pysim.TraceSpec(
        [p.pod_id for p in pods], [p.cpu_milli for p in pods], [p.memory_mib for p in pods],
        [p.num_gpu for p in pods], [p.gpu_milli for p in pods], [p.creation_time for p in pods],
        [p.duration_time for p in pods], [p.gpu_spec for p in pods])


def save_cluster(name, spec: pysim.ClusterSpec):
    n = spec.n_nodes
    ng = np.array([len(g) for g in spec.gpus], dtype=np.int32)
    gmem = np.zeros((n, 8), dtype=np.int32)
    gmil = np.zeros((n, 8), dtype=np.int32)
    for i, gl in enumerate(spec.gpus):
        for j, (m, t) in enumerate(gl):
            gmem[i, j], gmil[i, j] = m, t
    np.savez_compressed(os.path.join(DATA, f"cluster_{name}.npz"),
                        node_ids=np.array(spec.node_ids), cpu_total=np.array(spec.cpu_total, dtype=np.int32),
                        mem_total=np.array(spec.mem_total, dtype=np.int32), num_gpus=ng,
                        gpu_mem_total=gmem, gpu_milli_total=gmil)


def save_trace(name, spec: pysim.TraceSpec):
    np.savez_compressed(os.path.join(DATA, f"trace_{name}.npz"),
                        pod_ids=np.array(spec.pod_ids),
                        cpu_milli=np.array(spec.cpu_milli, dtype=np.int32),
                        memory_mib=np.array(spec.memory_mib, dtype=np.int32),
                        num_gpu=np.array(spec.num_gpu, dtype=np.int32),
                        gpu_milli=np.array(spec.gpu_milli, dtype=np.int32),
                        creation_time=np.array(spec.creation_time, dtype=np.int64),
                        duration_time=np.array(spec.duration_time, dtype=np.int64),
                        gpu_spec=np.array(spec.gpu_spec))


# ----------------------------------------------------------------------------- policies
def shipped_policies() -> dict:
    """name -> policy source defining priority_function(pod, node)."""
    out = {}
    for fname in ("first_fit_scheduler", "best_fit_scheduler", "funsearch_4901_scheduler",
                  "funsearch_4816_scheduler", "funsearch_4800_scheduler"):
        src = inspect.getsource(getattr(ref_ts, fname))
        src = re.sub(r"^def \w+\(pod: Pod, node: Node\) -> int:", "def priority_function(pod, node):", src, count=1)
        assert src.startswith("def priority_function(pod, node):"), src[:80]
        out["ts_" + fname.replace("_scheduler", "")] = src
    fs = SimpleFunSearch.__new__(SimpleFunSearch)
    for meth in ("_create_first_fit_policy", "_create_best_fit_policy", "_create_worst_fit_policy",
                 "_create_gpu_aware_policy", "_create_utilization_based_policy"):
        out["fs_" + meth[len("_create_"):-len("_policy")]] = getattr(fs, meth)()
    random.seed(20261106)
    for k in range(2):
        out[f"fs_random{k}"] = fs._create_random_policy()
    out["template_cpu_mem"] = PolicyTemplate.fill_template(
        "score = 1000.0 - (node.cpu_milli_left - pod.cpu_milli) / node.cpu_milli_total * 500.0\n"
        "    score -= (node.memory_mib_left - pod.memory_mib) / node.memory_mib_total * 300.0\n"
        "    if pod.num_gpu > 0:\n"
        "        score += 200.0 * (1 - (node.gpu_left - pod.num_gpu) / max(1, len(node.gpus)))")
    # the integer best-fit of the reference's tests/test_simulator.py:14-37
    ref_sim = _load_ref_module("ref_test_simulator", os.path.join(REF, "tests", "test_simulator.py"))
    src = inspect.getsource(ref_sim.best_fit_scheduler)
    src = re.sub(r"^def \w+\(pod: Pod, node: Node\) -> int:", "def priority_function(pod, node):", src, count=1)
    out["sim_int_best_fit"] = src
    return out


# ----------------------------------------------------------------------------- reference run
def run_reference(cluster, pods, scorer):
    """Runs the unmodified reference simulator; returns a dict in the oracle's vocabulary."""
    cluster = copy.deepcopy(cluster)
    pods = copy.deepcopy(pods)
    index = {id(p): i for i, p in enumerate(pods)}
    node_idx = {nid: i for i, nid in enumerate(cluster.nodes_dict)}
    des = DiscreteEventSimulator(pods)
    ev = SchedulingEvaluator(cluster, enabled=True)
    sim = KubernetesSimulator(cluster, pods, des, scorer, evaluator=ev)
    order = []
    real_push = des.push_deletion_event

    def logging_push(pod):
        order.append(index[id(pod)])
        real_push(pod)
    des.push_deletion_event = logging_push
    calls = [0]

    def counting(pod, node):
        calls[0] += 1
        return scorer(pod, node)
    sim.scheduler = counting
    err = ""
    t0 = time.perf_counter()
    try:
        sim.run_schedule()
    except Exception as exc:  # noqa: BLE001
        err = f"{type(exc).__name__}: {exc}"
    wall = time.perf_counter() - t0
    r = sim.get_evaluation_results()
    out = dict(error=err, wall=wall, events=ev.events_processed, scorer_calls=calls[0],
               placed=len(order), order=order, max_nodes=sim.max_nodes,
               final_time=[p.creation_time for p in pods],
               node_index=[node_idx.get(p.assigned_node, -1) for p in pods],
               gpu_lists=[list(p.assigned_gpus) for p in pods],
               num_snapshots=len(ev.utilization_snapshots), num_frag=len(ev.fragmentation_events),
               has_results=r is not None,
               score=float(ev.get_policy_score(pods)) if not err else 0.0)
    if r is not None:
        out.update(avg_cpu=r.avg_cpu_utilization, avg_mem=r.avg_memory_utilization,
                   avg_gpu_count=r.avg_gpu_count_utilization, avg_gpu_milli=r.avg_gpu_memory_utilization,
                   fragmentation=r.gpu_fragmentation_score)
    out["cpu_left"] = [n.cpu_milli_left for n in cluster.nodes_dict.values()]
    out["mem_left"] = [n.memory_mib_left for n in cluster.nodes_dict.values()]
    out["gpu_left"] = [n.gpu_left for n in cluster.nodes_dict.values()]
    out["gpu_milli_left"] = [[g.gpu_milli_left for g in n.gpus] for n in cluster.nodes_dict.values()]
    return out


def check_equal(tag, ref, orc: pysim.EpisodeResult):
    def eq(name, a, b):
        assert a == b, f"{tag}: {name} differs: reference {a!r} vs oracle {b!r}"
    eq("error?", bool(ref["error"]), orc.status != 0)
    eq("events", ref["events"], orc.events)
    eq("scorer_calls", ref["scorer_calls"], orc.scorer_calls)
    eq("placed", ref["placed"], orc.placed)
    eq("order", ref["order"], orc.placement_order)
    eq("final_time", ref["final_time"], orc.final_time)
    eq("node_index", ref["node_index"], orc.node_index)
    eq("gpu_lists", ref["gpu_lists"], orc.gpu_lists)
    eq("max_nodes", ref["max_nodes"], orc.max_nodes)
    eq("num_snapshots", ref["num_snapshots"], orc.num_snapshots)
    eq("num_frag", ref["num_frag"], orc.num_fragmentation_events)
    eq("has_results", ref["has_results"], orc.has_results)
    if ref["has_results"]:
        for k, v in (("avg_cpu", orc.avg_cpu), ("avg_mem", orc.avg_mem), ("avg_gpu_count", orc.avg_gpu_count),
                     ("avg_gpu_milli", orc.avg_gpu_milli), ("fragmentation", orc.fragmentation)):
            eq(k, ref[k], v)
    eq("score", ref["score"], orc.score)
    for k in ("cpu_left", "mem_left", "gpu_left", "gpu_milli_left"):
        eq(k, ref[k], getattr(orc, k))


def golden_record(trace_spec, cluster_spec, orc: pysim.EpisodeResult, ref_wall):
    return dict(status=orc.status, error=orc.error, events=orc.events, attempts=orc.attempts,
                scorer_calls=orc.scorer_calls, placed=orc.placed, num_snapshots=orc.num_snapshots,
                num_fragmentation_events=orc.num_fragmentation_events, max_nodes=orc.max_nodes,
                max_waiting=orc.max_waiting, last_event_time=orc.last_event_time,
                avg_cpu=orc.avg_cpu.hex(), avg_mem=orc.avg_mem.hex(), avg_gpu_count=orc.avg_gpu_count.hex(),
                avg_gpu_milli=orc.avg_gpu_milli.hex(), fragmentation=orc.fragmentation.hex(),
                has_results=orc.has_results, score=float(orc.score).hex(), score_repr=repr(float(orc.score)),
                digest=pysim.placement_digest(trace_spec, cluster_spec, orc),
                digest_alloc_order=pysim.placement_digest_allocation_order(trace_spec, cluster_spec, orc),
                state_digest=hashlib.sha256(json.dumps(
                    [orc.final_time, orc.node_index, orc.gpu_mask, orc.cpu_left, orc.mem_left, orc.gpu_left,
                     orc.gpu_milli_left]).encode()).hexdigest(),
                reference_wall_s=round(ref_wall, 4))


# ----------------------------------------------------------------------------- edge cases
def edge_cases():
    """Small hand-built workloads probing the corner semantics of SURVEY.md A.9."""
    def cl(rows):
        return pysim.ClusterSpec([f"n{i}" for i in range(len(rows))], [r[0] for r in rows], [r[1] for r in rows],
                                 [[(16000, 1000)] * r[2] for r in rows])

    def tr(rows, ids=None):
        ids = ids or [f"p{i}" for i in range(len(rows))]
        return pysim.TraceSpec(ids, [r[0] for r in rows], [r[1] for r in rows], [r[2] for r in rows],
                               [r[3] for r in rows], [r[4] for r in rows], [r[5] for r in rows], [""] * len(rows))
    cases = {}
    # the reference's tests/test_simulator.py scenario (2 nodes, 4 pods)
    cases["test_simulator"] = (
        pysim.ClusterSpec(["node1", "node2"], [8000, 4000], [16000, 8000], [[(8000, 1000)] * 2, []]),
        pysim.TraceSpec(["pod1", "pod2", "pod3", "pod4"], [1000, 2000, 3000, 1500], [2000, 4000, 6000, 3000],
                        [0, 1, 0, 2], [0, 500, 0, 400], [0, 5, 10, 15], [10, 15, 8, 12], ["", "v100", "", "v100"]),
        "sim_int_best_fit")
    # (1) oversized pod, nothing pending -> dropped
    cases["oversized_dropped"] = (cl([(1000, 1000, 0)]), tr([(5000, 10, 0, 0, 3, 5)]), "fs_first_fit")
    # (2) oversized pod retried after each deletion, then dropped
    cases["oversized_retried"] = (cl([(4000, 4000, 1)]),
                                  tr([(1000, 100, 0, 0, 0, 5), (1000, 100, 0, 0, 2, 50), (9000, 10, 0, 0, 1, 7)]),
                                  "fs_first_fit")
    # (6) same-time deletion/creation ordered by pod id, both id orders
    cases["same_time_del_first"] = (cl([(1000, 1000, 0)]), tr([(1000, 10, 0, 0, 0, 10), (1000, 10, 0, 0, 10, 5)],
                                                               ids=["p0", "p1"]), "fs_first_fit")
    cases["same_time_create_first"] = (cl([(1000, 1000, 0)]), tr([(1000, 10, 0, 0, 0, 10), (1000, 10, 0, 0, 10, 5)],
                                                                  ids=["p1", "p0"]), "fs_first_fit")
    # (7) single zero-duration pod
    cases["single_zero_duration"] = (cl([(64000, 262144, 2)]), tr([(12000, 16384, 1, 1000, 0, 0)]), "fs_first_fit")
    # fractional GPU sharing and multi-GPU best fit
    cases["gpu_sharing"] = (cl([(64000, 262144, 4), (64000, 262144, 2)]),
                            tr([(1000, 100, 1, 300, 0, 100), (1000, 100, 1, 800, 1, 100), (1000, 100, 2, 1000, 2, 50),
                                (1000, 100, 1, 200, 3, 10), (1000, 100, 1, 700, 4, 100), (1000, 100, 2, 1000, 5, 100),
                                (1000, 100, 1, 650, 6, 30), (1000, 100, 4, 1000, 7, 10), (500, 50, 0, 0, 8, 1)]),
                            "ts_best_fit")
    # unsorted pod list with mixed-width ids (string order != numeric order)
    cases["unsorted_mixed_ids"] = (cl([(8000, 8000, 1), (8000, 8000, 0)]),
                                   tr([(4000, 100, 0, 0, 50, 10), (4000, 100, 1, 500, 5, 100), (4000, 100, 0, 0, 5, 20),
                                       (6000, 100, 0, 0, 7, 30), (4000, 100, 1, 600, 5, 15), (100, 100, 0, 0, 0, 1)],
                                      ids=["pod-10", "pod-9", "pod-100", "pod-1", "pod-99", "pod-2"]),
                                   "ts_funsearch_4816")
    return cases


RAISING_POLICIES = {
    # (3) scorer ignoring gpu_milli -> ValueError("Not enough GPUs") inside the allocator
    "x_ignores_gpu_milli": "def priority_function(pod, node):\n"
                           "    if pod.cpu_milli > node.cpu_milli_left or pod.num_gpu > node.gpu_left:\n"
                           "        return 0\n    return 10\n",
    # (4) scorer ignoring capacity -> residuals go negative, no error
    "x_ignores_capacity": "def priority_function(pod, node):\n    return 5\n",
    # ZeroDivisionError on CPU-only nodes (tests/test_integration.py:42 style)
    "x_div_by_len_gpus": "def priority_function(pod, node):\n"
                         "    if pod.cpu_milli > node.cpu_milli_left or pod.memory_mib > node.memory_mib_left or "
                         "pod.num_gpu > node.gpu_left:\n        return 0\n"
                         "    return 1 + int(100 * (node.gpu_left - pod.num_gpu) / len(node.gpus))\n",
}


def main():
    os.makedirs(DATA, exist_ok=True)
    parser = TraceParser()
    execu = SafeExecutor(3)

    clusters = {}
    for cname, cfile in (("gpu16", "gpu_models_filtered.csv"), ("all1523", "openb_node_list_all_node.csv"),
                         ("gpu1213", "openb_node_list_gpu_node.csv")):
        clusters[cname] = parser.parse_cluster(cfile)
        save_cluster(cname, cluster_to_spec(clusters[cname]))

    traces = {}
    for pf in parser.get_available_pod_files():
        name = pf[len("openb_pod_list_"):-len(".csv")]
        try:
            traces[name] = parser.parse_pods(pf)
        except KeyError:
            print(f"skip {pf}: no time columns")
            continue
        save_trace(name, pods_to_spec(traces[name]))

    policies = shipped_policies()
    for name, src in policies.items():
        execu.validate_code_content(src) if not name.startswith(("ts_", "sim_")) else None
        execu.validate_code_structure(src)
    allp = dict(policies)
    allp.update(RAISING_POLICIES)
    with open(os.path.join(GOLD, "policies.json"), "w") as f:
        json.dump(allp, f, indent=1, sort_keys=True)

    runs = {}
    meta = dict(python=sys.version, libc=" ".join(platform.libc_ver()), machine=platform.machine(),
                generated_by="tests/golden/make_golden.py", reference="ttanv/funsearch-kubernetes-simulator")

    def one(tag, cluster, pods, cspec, tspec, pname):
        src = allp[pname]
        ref_scorer = FunSearchScheduler(src, execu)
        ref_scorer.fallback_print = False
        import io
        import contextlib
        with contextlib.redirect_stdout(io.StringIO()):
            ref = run_reference(cluster, pods, ref_scorer)
            orc = pysim.run_episode(cspec, tspec, pysim.scorer_from_source(src))
        check_equal(tag, ref, orc)
        runs[tag] = golden_record(tspec, cspec, orc, ref["wall"])
        print(f"ok {tag:44s} ev={orc.events:6d} att={orc.attempts:6d} snap={orc.num_snapshots:3d} "
              f"frag={orc.num_fragmentation_events:6d} st={orc.status} score={orc.score!r}")

    c16 = clusters["gpu16"]
    c16s = cluster_to_spec(c16)
    main_policies = [p for p in policies if p != "sim_int_best_fit"]
    for tname, pods in traces.items():
        tspec = pods_to_spec(pods)
        for pname in main_policies:
            one(f"{tname}|gpu16|{pname}", c16, pods, c16s, tspec, pname)
    for pname in RAISING_POLICIES:
        one(f"default|gpu16|{pname}", c16, traces["default"], c16s, pods_to_spec(traces["default"]), pname)

    # default trace on the first 1,024 rows of the all-node list (config-3 cluster shape)
    big = clusters["all1523"]
    big1024 = Cluster(nodes_dict=dict(list(big.nodes_dict.items())[:1024]))
    b_s = cluster_to_spec(big1024)
    for pname in ("ts_first_fit", "ts_best_fit", "ts_funsearch_4816"):
        one(f"default|all1024|{pname}", big1024, traces["default"], b_s, pods_to_spec(traces["default"]), pname)
    # a 64-node slice with a 2,000-pod prefix of a busy trace (multi-node-per-lane shape with retries)
    c64 = Cluster(nodes_dict=dict(list(clusters["gpu1213"].nodes_dict.items())[:64]))
    pods2k = traces["cpu300"][:4000]
    for pname in ("ts_best_fit", "ts_funsearch_4901", "ts_funsearch_4800"):
        one(f"cpu300[:4000]|gpu64|{pname}", c64, pods2k, cluster_to_spec(c64), pods_to_spec(pods2k), pname)

    # hand-built edge cases
    edge = {}
    for ename, (cspec, tspec, pname) in edge_cases().items():
        nodes = {}
        for i, nid in enumerate(cspec.node_ids):
            gl = [GPU(m, m, t, t) for (m, t) in cspec.gpus[i]]
            nodes[nid] = Node(nid, cspec.cpu_total[i], cspec.cpu_total[i], cspec.mem_total[i], cspec.mem_total[i],
                              len(gl), gl)
        pods = [Pod(tspec.pod_ids[i], tspec.cpu_milli[i], tspec.memory_mib[i], tspec.num_gpu[i], tspec.gpu_milli[i],
                    tspec.gpu_spec[i], tspec.creation_time[i], tspec.duration_time[i], "", [])
                for i in range(tspec.n_pods)]
        one(f"edge:{ename}|{pname}", Cluster(nodes_dict=nodes), pods, cspec, tspec, pname)
        edge[ename] = dict(policy=pname,
                           cluster=dict(node_ids=cspec.node_ids, cpu_total=cspec.cpu_total, mem_total=cspec.mem_total,
                                        gpus=cspec.gpus),
                           trace=dict(pod_ids=tspec.pod_ids, cpu_milli=tspec.cpu_milli, memory_mib=tspec.memory_mib,
                                      num_gpu=tspec.num_gpu, gpu_milli=tspec.gpu_milli,
                                      creation_time=tspec.creation_time, duration_time=tspec.duration_time))

    with open(os.path.join(GOLD, "golden_runs.json"), "w") as f:
        json.dump(dict(meta=meta, runs=runs, edge_cases=edge), f, indent=1, sort_keys=True)
    print(f"wrote {len(runs)} golden runs")


if __name__ == "__main__":
    main()
