"""CPU oracle (TEST INFRASTRUCTURE, never the product path): a flat-array Python
restatement of the reference's episode replay.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; it is the checker, not the thing measured or shipped.

Parity status: PINNED.  tests/golden/make_golden.py runs this file side by side with
the real reference (imported from /root/reference in the build container) on every
bundled trace x every shipped policy and asserts identical placement logs, counters
and metrics before it writes the committed fixtures under tests/golden/.

What it restates (reference file:line):
  * event engine       simulator/event_simulator.py:19-59   (heapq of (time, Event);
                       Event order = pod_id string order, :16-17; retry = first
                       DELETION in heap *array* order, :51-59)
  * event loop         simulator/main.py:50-72
  * creation/deletion  simulator/main.py:74-148
  * GPU best fit       simulator/main.py:150-177
  * evaluator          simulator/evaluator.py:27-164
  * policy wrapper     funsearch/funsearch_integration.py:67-101 (int(max(0, score)))
  * policy sandbox env funsearch/safe_execution.py:98-124

The restatement keeps CPython's own ``heapq`` (so the heap *layout* semantics the retry
scan depends on are the interpreter's, exactly as in the reference) but replaces the
object graph by flat per-pod / per-node lists where the simulator itself is concerned;
policies still see attribute-style ``pod`` / ``node`` / ``gpu`` views because that is
the plugin contract.
"""
from __future__ import annotations

import heapq
import math
import operator
import statistics
from dataclasses import dataclass, field
from typing import Callable, Optional, Sequence

# status words shared with the CUDA path (include/ks_b200.h)
ST_OK = 0
ST_POLICY_EXCEPTION = 1
ST_GPU_SHORTFALL = 2
ST_STEP_BUDGET = 3
ST_UNSUPPORTED = 4

_CREATE = 0
_DELETE = 1


class GpuView:
    __slots__ = ("memory_mib_left", "memory_mib_total", "gpu_milli_left", "gpu_milli_total")

    def __init__(self, mem_total: int, milli_total: int):
        self.memory_mib_left = mem_total
        self.memory_mib_total = mem_total
        self.gpu_milli_left = milli_total
        self.gpu_milli_total = milli_total


class NodeView:
    __slots__ = ("node_id", "cpu_milli_left", "cpu_milli_total", "memory_mib_left",
                 "memory_mib_total", "gpu_left", "gpus")


class PodView:
    __slots__ = ("pod_id", "cpu_milli", "memory_mib", "num_gpu", "gpu_milli", "gpu_spec",
                 "creation_time", "duration_time", "assigned_node", "assigned_gpus", "index")


@dataclass
class ClusterSpec:
    """Immutable cluster template: one row per node, in nodes_dict insertion order."""
    node_ids: list
    cpu_total: list
    mem_total: list
    # per node: list of (gpu_memory_mib_total, gpu_milli_total)
    gpus: list

    @property
    def n_nodes(self) -> int:
        return len(self.node_ids)


@dataclass
class TraceSpec:
    """Immutable pod trace template in pod_list order."""
    pod_ids: list
    cpu_milli: list
    memory_mib: list
    num_gpu: list
    gpu_milli: list
    creation_time: list
    duration_time: list
    gpu_spec: Optional[list] = None

    @property
    def n_pods(self) -> int:
        return len(self.pod_ids)


@dataclass
class EpisodeResult:
    status: int = ST_OK
    error: str = ""
    events: int = 0
    attempts: int = 0
    scorer_calls: int = 0
    placed: int = 0
    num_snapshots: int = 0
    num_fragmentation_events: int = 0
    max_nodes: int = 0
    max_waiting: int = 0
    last_event_time: int = 0
    avg_cpu: float = 0.0
    avg_mem: float = 0.0
    avg_gpu_count: float = 0.0
    avg_gpu_milli: float = 0.0
    fragmentation: float = 0.0
    has_results: bool = False
    score: float = 0.0
    # per pod (pod_list order): final creation time, node index (-1 = never placed), GPU bitmask
    final_time: list = field(default_factory=list)
    node_index: list = field(default_factory=list)
    gpu_mask: list = field(default_factory=list)
    # per pod: GPU indices in the allocator's own (best-fit) order
    gpu_lists: list = field(default_factory=list)
    # pods in order of successful placement
    placement_order: list = field(default_factory=list)
    # final residuals, for accounting checks
    cpu_left: list = field(default_factory=list)
    mem_left: list = field(default_factory=list)
    gpu_left: list = field(default_factory=list)
    gpu_milli_left: list = field(default_factory=list)


def sandbox_env() -> dict:
    """The restricted globals a policy source is exec'd in (safe_execution.py:98-124)."""
    import builtins
    names = ("abs", "min", "max", "sum", "len", "range", "enumerate",
             "int", "float", "bool", "str", "round", "sorted")
    env = {"__builtins__": {n: getattr(builtins, n) for n in names}}
    env["math"] = type("M", (), {n: staticmethod(getattr(math, n))
                                 for n in ("sqrt", "log", "exp", "pow", "sin", "cos", "tan")})()
    env["operator"] = type("O", (), {n: staticmethod(getattr(operator, n))
                                     for n in ("add", "sub", "mul", "truediv", "mod")})()
    return env


def scorer_from_source(code: str) -> Callable:
    """Policy source -> wrapped scorer, as FunSearchScheduler does
    (funsearch_integration.py:77-101): exec once, then int(max(0, f(pod, node)))."""
    env = sandbox_env()
    exec(code, env)
    fn = env.get("priority_function")
    if not fn:
        raise ValueError("No priority_function found in evolved code")

    def wrapped(pod, node):
        return int(max(0, fn(pod, node)))
    return wrapped


def _build_nodes(cluster: ClusterSpec) -> list:
    nodes = []
    for i in range(cluster.n_nodes):
        nv = NodeView()
        nv.node_id = cluster.node_ids[i]
        nv.cpu_milli_left = nv.cpu_milli_total = cluster.cpu_total[i]
        nv.memory_mib_left = nv.memory_mib_total = cluster.mem_total[i]
        nv.gpus = [GpuView(m, t) for (m, t) in cluster.gpus[i]]
        nv.gpu_left = len(nv.gpus)
        nodes.append(nv)
    return nodes


def _build_pods(trace: TraceSpec) -> list:
    pods = []
    spec = trace.gpu_spec
    for i in range(trace.n_pods):
        pv = PodView()
        pv.index = i
        pv.pod_id = trace.pod_ids[i]
        pv.cpu_milli = trace.cpu_milli[i]
        pv.memory_mib = trace.memory_mib[i]
        pv.num_gpu = trace.num_gpu[i]
        pv.gpu_milli = trace.gpu_milli[i]
        pv.gpu_spec = spec[i] if spec is not None else ""
        pv.creation_time = trace.creation_time[i]
        pv.duration_time = trace.duration_time[i]
        pv.assigned_node = ""
        pv.assigned_gpus = []
        pods.append(pv)
    return pods


def run_episode(cluster: ClusterSpec, trace: TraceSpec, scorer: Callable,
                snapshot_interval: float = 0.05,
                max_events: Optional[int] = None) -> EpisodeResult:
    """Replay ``trace`` on a fresh copy of ``cluster`` under ``scorer(pod, node) -> int``.

    Mirrors KubernetesSimulator.run_schedule with a SchedulingEvaluator attached
    (main.py:50-72, evaluator.py).  Exceptions raised by the scorer or by the GPU
    allocator end the episode and are reported through ``status`` (the reference's
    worker turns them into fitness 0, funsearch_integration.py:63-64).
    """
    res = EpisodeResult()
    nodes = _build_nodes(cluster)
    pods = _build_pods(trace)
    n_pods = len(pods)
    node_index = {nv.node_id: i for i, nv in enumerate(nodes)}

    # evaluator constants (evaluator.py:35-38)
    tot_cpu = sum(nv.cpu_milli_total for nv in nodes)
    tot_mem = sum(nv.memory_mib_total for nv in nodes)
    tot_gcount = sum(len(nv.gpus) for nv in nodes)
    tot_gmilli = sum(g.gpu_milli_total for nv in nodes for g in nv.gpus)

    # heap items: (time, pod_id, kind, pod).  pod ids are unique, so tuple order is decided by
    # (time, pod_id) exactly as (time, Event) with Event.__lt__ on pod_id (event_simulator.py:16-17).
    heap = [(p.creation_time, p.pod_id, _CREATE, p) for p in pods]
    heapq.heapify(heap)
    total_events = len(heap)

    processed = 0
    threshold = snapshot_interval
    s_cpu, s_mem, s_gc, s_gm = [], [], [], []
    frag = []
    waiting = []          # pods that failed at least once and are not placed (main.py:42)
    max_nodes = 0
    res.final_time = [p.creation_time for p in pods]
    res.node_index = [-1] * n_pods
    res.gpu_mask = [0] * n_pods
    res.gpu_lists = [[] for _ in range(n_pods)]

    try:
        while heap:
            if max_events is not None and processed >= max_events:
                res.status = ST_STEP_BUDGET
                res.error = "step budget exhausted"
                break
            t, _pid, kind, pod = heapq.heappop(heap)
            res.last_event_time = t
            if kind == _DELETE:
                # main.py:74-99
                nv = nodes[node_index[pod.assigned_node]]
                nv.cpu_milli_left += pod.cpu_milli
                nv.memory_mib_left += pod.memory_mib
                nv.gpu_left += pod.num_gpu
                for gi in pod.assigned_gpus:
                    nv.gpus[gi].gpu_milli_left += pod.gpu_milli
            else:
                # main.py:101-148
                res.attempts += 1
                best, best_nv = 0, None
                for nv in nodes:
                    res.scorer_calls += 1
                    s = scorer(pod, nv)
                    if s > best:
                        best, best_nv = s, nv
                if best_nv is None:
                    if not any(w is pod for w in waiting):
                        waiting.append(pod)
                        res.max_waiting = max(res.max_waiting, len(waiting))
                    # evaluator.py:69-75,144-164
                    need = [w.gpu_milli for w in waiting if w.num_gpu > 0]
                    if need:
                        m = min(need)
                        lost = 0
                        for nv in nodes:
                            for g in nv.gpus:
                                if 0 < g.gpu_milli_left < m:
                                    lost += g.gpu_milli_left
                        frag.append(lost / tot_gmilli if tot_gmilli > 0 else 0.0)
                    else:
                        frag.append(0.0)
                    # event_simulator.py:51-59: first DELETION in array order
                    for (ht, _hp, hk, _hpod) in heap:
                        if hk == _DELETE:
                            pod.creation_time = ht + 1
                            heapq.heappush(heap, (ht + 1, pod.pod_id, _CREATE, pod))
                            break
                else:
                    best_nv.cpu_milli_left -= pod.cpu_milli
                    best_nv.memory_mib_left -= pod.memory_mib
                    best_nv.gpu_left -= pod.num_gpu
                    chosen = []
                    if pod.num_gpu != 0:
                        # main.py:150-177
                        cand = [(gi, g.gpu_milli_left) for gi, g in enumerate(best_nv.gpus)
                                if g.gpu_milli_left >= pod.gpu_milli]
                        if len(cand) < pod.num_gpu:
                            res.status = ST_GPU_SHORTFALL
                            raise ValueError(f"Not enough GPUs available on node {best_nv.node_id}")
                        cand.sort(key=lambda c: c[1])
                        for k in range(pod.num_gpu):
                            gi = cand[k][0]
                            chosen.append(gi)
                            best_nv.gpus[gi].gpu_milli_left -= pod.gpu_milli
                    pod.assigned_node = best_nv.node_id
                    pod.assigned_gpus = chosen
                    for wi, w in enumerate(waiting):
                        if w is pod:
                            del waiting[wi]
                            break
                    heapq.heappush(heap, (pod.creation_time + pod.duration_time, pod.pod_id, _DELETE, pod))
                    res.placed += 1
                    res.placement_order.append(pod.index)
                    res.final_time[pod.index] = pod.creation_time
                    res.node_index[pod.index] = node_index[best_nv.node_id]
                    mask = 0
                    for gi in chosen:
                        mask |= 1 << gi
                    res.gpu_mask[pod.index] = mask
                    res.gpu_lists[pod.index] = chosen

            # evaluator.py:55-67 + 129-142
            processed += 1
            progress = processed / total_events if total_events > 0 else 0
            if progress >= threshold:
                u_cpu = sum(nv.cpu_milli_total - nv.cpu_milli_left for nv in nodes)
                u_mem = sum(nv.memory_mib_total - nv.memory_mib_left for nv in nodes)
                u_gc = sum(len(nv.gpus) - nv.gpu_left for nv in nodes)
                u_gm = sum(g.gpu_milli_total - g.gpu_milli_left for nv in nodes for g in nv.gpus)
                s_cpu.append(u_cpu / tot_cpu if tot_cpu > 0 else 0.0)
                s_mem.append(u_mem / tot_mem if tot_mem > 0 else 0.0)
                s_gc.append(u_gc / tot_gcount if tot_gcount > 0 else 0.0)
                s_gm.append(u_gm / tot_gmilli if tot_gmilli > 0 else 0.0)
                threshold += snapshot_interval
            # main.py:68-72
            active = 0
            for nv in nodes:
                if (nv.cpu_milli_left < nv.cpu_milli_total or nv.memory_mib_left < nv.memory_mib_total
                        or nv.gpu_left < len(nv.gpus)):
                    active += 1
            if active > max_nodes:
                max_nodes = active
    except Exception as exc:  # noqa: BLE001 - the reference's worker catches everything
        if res.status == ST_OK:
            res.status = ST_POLICY_EXCEPTION
        res.error = f"{type(exc).__name__}: {exc}"

    res.events = processed
    res.max_nodes = max_nodes
    res.num_snapshots = len(s_cpu)
    res.num_fragmentation_events = len(frag)
    res.cpu_left = [nv.cpu_milli_left for nv in nodes]
    res.mem_left = [nv.memory_mib_left for nv in nodes]
    res.gpu_left = [nv.gpu_left for nv in nodes]
    res.gpu_milli_left = [[g.gpu_milli_left for g in nv.gpus] for nv in nodes]
    for p in pods:
        res.final_time[p.index] = p.creation_time
    if s_cpu:
        # evaluator.py:77-99
        res.has_results = True
        res.avg_cpu = statistics.mean(s_cpu)
        res.avg_mem = statistics.mean(s_mem)
        res.avg_gpu_count = statistics.mean(s_gc)
        res.avg_gpu_milli = statistics.mean(s_gm)
        res.fragmentation = statistics.mean(frag) if frag else 0.0
    # evaluator.py:101-127 (an aborted episode scores 0: funsearch_integration.py:63-64)
    if res.status != ST_OK or not res.has_results:
        res.score = 0.0
    elif any(ni < 0 for ni in res.node_index):
        res.score = 0.0
    else:
        overall = (res.avg_cpu + res.avg_mem + res.avg_gpu_count + res.avg_gpu_milli) / 4.0
        res.score = max(0.0, min(1.0, overall - min(0.1, res.fragmentation)))
    return res


def placement_digest(trace: TraceSpec, cluster: ClusterSpec, res: EpisodeResult) -> str:
    """SHA-256 over 'pod_id,final_creation_time,node_id,g0|g1|...' lines in order of
    successful placement (the digest format of SURVEY.md App. D)."""
    import hashlib
    lines = []
    for pi in res.placement_order:
        m = res.gpu_mask[pi]
        gl = "|".join(str(g) for g in range(32) if (m >> g) & 1)
        lines.append(f"{trace.pod_ids[pi]},{res.final_time[pi]},{cluster.node_ids[res.node_index[pi]]},{gl}")
    return hashlib.sha256("\n".join(lines).encode()).hexdigest()


def placement_digest_allocation_order(trace: TraceSpec, cluster: ClusterSpec, res: EpisodeResult) -> str:
    """Same digest, but with the GPU indices in the allocator's own (best-fit) order: this is
    the byte format whose first 16 hex digits SURVEY.md App. D lists."""
    import hashlib
    lines = []
    for pi in res.placement_order:
        gl = "|".join(str(g) for g in res.gpu_lists[pi])
        lines.append(f"{trace.pod_ids[pi]},{res.final_time[pi]},{cluster.node_ids[res.node_index[pi]]},{gl}")
    return hashlib.sha256("\n".join(lines).encode()).hexdigest()
