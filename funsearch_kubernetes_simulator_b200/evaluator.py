"""Batched fitness evaluation on a B200: the drop-in for the reference's evaluation call sites.

Reference boundary:
  * evaluate_policy_standalone((idx, code)) -> (idx, code, score)   funsearch/funsearch_integration.py:30-64
  * SimpleFunSearch._evaluate_policy_full(code) -> Optional[float]    funsearch/funsearch_integration.py:433-459
  * the ProcessPoolExecutor block that maps the former over a generation          :535-540
  * EvaluationResults / get_policy_score                               simulator/evaluator.py:16-25, 101-127

`BatchEvaluator` holds traces and one cluster resident in HBM and evaluates any number of
(policy, trace) episodes with ONE launch of the replay kernel.  PyTorch is used for device
buffers, streams and (multi-GPU) torch.distributed only; every episode runs in the CUDA kernel
of libks_b200.so.  No CPU path exists: without the library or without a GPU the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np

from . import _native as nat
from .policy_compiler import CompiledPolicy, PolicyCompileError, compile_policy
from .workload import ClusterSpec, Trace


@dataclass
class EvaluationResults:
    """Same fields as the reference's EvaluationResults (simulator/evaluator.py:16-25)."""
    avg_cpu_utilization: float
    avg_memory_utilization: float
    avg_gpu_count_utilization: float
    avg_gpu_memory_utilization: float
    gpu_fragmentation_score: float
    num_snapshots: int
    num_fragmentation_events: int


@dataclass
class EpisodeOutcome:
    """One replayed episode."""
    status: int
    policy_score: float                      # get_policy_score(); 0 for a failed episode (:63-64)
    results: Optional[EvaluationResults]     # None when no snapshot was taken (evaluator.py:79-80)
    events: int
    attempts: int
    scorer_calls: int
    placed: int
    max_nodes: int
    max_waiting: int
    last_event_time: int
    device_cycles: int
    # only with want_placements: per pod, in pod_list order
    final_creation_time: Optional[np.ndarray] = None
    node_index: Optional[np.ndarray] = None   # -1 = never placed
    gpu_mask: Optional[np.ndarray] = None
    waiting: Optional[np.ndarray] = None      # bool: member of waiting_pods when the replay ended (main.py:116-117)

    @property
    def status_name(self) -> str:
        return nat.STATUS_NAMES.get(self.status, str(self.status))

    def placement_order(self, id_rank: np.ndarray) -> np.ndarray:
        """Pods in order of successful placement.  Events pop in non-decreasing (time, id_rank)
        order and a placement happens at its pod's final creation time, so sorting placed pods by
        (final time, id_rank) reproduces the order."""
        placed = np.nonzero(self.node_index >= 0)[0]
        key = np.lexsort((id_rank[placed], self.final_creation_time[placed]))
        return placed[key]


def table_order(ep_policy: np.ndarray, ep_const: Optional[np.ndarray] = None) -> Optional[np.ndarray]:
    """Stable permutation that makes the episode table structure-major (and policy-major inside a structure), or None
    when it already is.  The kernel hands each CTA a contiguous range of the table, so episodes that run the same
    device function sit side by side on one SM and share its code in the instruction cache
    (csrc/ks_sim_kernel.cu, work distribution)."""
    ep_policy = np.asarray(ep_policy)
    if ep_policy.size < 2:
        return None
    if ep_const is None:
        if np.all(ep_policy[1:] >= ep_policy[:-1]):
            return None
        return np.argsort(ep_policy, kind="stable")
    key = ep_policy.astype(np.int64) * (1 << 32) + np.asarray(ep_const).astype(np.int64)
    if np.all(key[1:] >= key[:-1]):
        return None
    return np.argsort(key, kind="stable")


@dataclass
class BoundPolicies:
    """A list of compiled policies bound to a linked program: the table slot (structure) and the start of the float
    literals of each, plus the literals of all of them back to back."""
    program: int
    slot: np.ndarray        # int32 [n_policies]
    const_off: np.ndarray   # int32 [n_policies]
    consts: np.ndarray      # float64


@dataclass
class PopulationReport:
    """What evaluate_population did with each source, next to the scores."""
    scores: List[float]
    status: List[int]            # worst per-episode status of the policy (0 = ok); KS_ST_UNSUPPORTED also for sources
                                 # the translator declined
    not_evaluated: List[int]     # indices whose fitness is NOT the reference's: outside the translated subset
    messages: dict               # index -> why (translator message or status name)
    distinct_structures: int = 0


class BatchEvaluator:
    """Traces + one cluster resident on one GPU; evaluates batches of episodes."""

    def __init__(self, cluster: ClusterSpec, traces: Sequence[Trace], device: int = 0,
                 snapshot_interval: float = 0.05, cache_dir: Optional[str] = None, processes: Optional[int] = None):
        import torch  # device buffers only
        self._torch = torch
        if not torch.cuda.is_available():
            raise nat.KsError(-2, "no CUDA device: the batched evaluator has no CPU fallback")
        self.device = torch.device("cuda", device)
        self.lib = nat.lib()
        nat.check(self.lib.ks_init(device))
        if cache_dir:
            nat.check(self.lib.ks_set_cache_dir(cache_dir.encode()))
        self.cluster = cluster
        self.traces = list(traces)
        self.snapshot_interval = snapshot_interval
        nconst, totals = cluster.packed()
        self._totals = totals
        self._nconst = torch.from_numpy(nconst).to(self.device)
        self._trace_bufs = []
        table = np.zeros((len(self.traces), 4), dtype=np.int64)
        for i, tr in enumerate(self.traces):
            _rank, podrec, heap0, place0 = tr.packed()
            d_pod = torch.from_numpy(podrec).to(self.device)
            d_heap = torch.from_numpy(heap0.view(np.int64)).to(self.device)
            d_place = torch.from_numpy(place0.view(np.int64)).to(self.device)
            self._trace_bufs.append((d_pod, d_heap, d_place))
            table[i] = (d_pod.data_ptr(), d_heap.data_ptr(), d_place.data_ptr(), tr.n_pods)
        self._trace_table = torch.from_numpy(table).to(self.device)
        self.max_pods = max([t.n_pods for t in self.traces] + [1])
        self._programs: List[Tuple[frozenset, dict, int]] = []   # (structure handles, handle -> slot, program), newest last
        self._pinned: set = set()
        self._scratch: dict = {}   # per stream
        self.last_plan = None
        self.last_kernel_ms = None
        self.last_compile_stats: dict = {}
        self.last_report: Optional[PopulationReport] = None
        self.sm_count = torch.cuda.get_device_properties(self.device).multi_processor_count
        self.warps_per_team = 0   # 0 = planner's choice; set to force a team width
        self.ctas_per_sm = 0      # 0 = planner's choice; 1 = latency kernel, 2 = dense (throughput) kernel
        # Compile workers are started now, not inside the first evaluation: they warm up (imports, libnvrtc) while the
        # caller is still generating its first population.
        if processes is None:
            # this process's share of the host: ranks of one torchrun launch compile their populations side by side
            try:
                cores = len(os.sched_getaffinity(0))
            except AttributeError:
                cores = os.cpu_count() or 1
            ranks_here = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1") or 1))
            processes = max(1, min(cores // ranks_here, 32))
        self.processes = processes
        if self.processes > 1 and os.environ.get("KS_NO_POOL") != "1":
            from . import compile_pool
            d = compile_pool.start(self.processes, self.COMPILE_BATCH)
            if not cache_dir:
                nat.check(self.lib.ks_set_cache_dir(d.encode()))
        nat.check(self.lib.ks_set_compile_batch(self.COMPILE_BATCH))

    # ------------------------------------------------------------------ policies
    COMPILE_BATCH = 16          # structures per NVRTC program (ks_set_compile_batch)
    POOL_SOURCES = 48           # fresh sources from which translating in the pool pays off
    POOL_STRUCTURES = 24        # new structures from which NVRTC in the pool pays off
    MAX_PROGRAMS = 12           # linked modules kept loaded (least recently used ones are released)

    def compile(self, sources: Sequence[str], threads: int = 0, processes: Optional[int] = None) -> List[Optional[CompiledPolicy]]:
        """Python policy sources -> compiled policies (structure function + float literals).
        A source outside the supported subset yields None (its episodes are reported as not evaluated).
        A generation of fresh sources is translated by the pool of worker processes started with the evaluator
        (`processes=0` keeps everything in this process); only structures never seen before reach NVRTC, in groups
        of COMPILE_BATCH, again on the pool."""
        import time
        from . import compile_pool, policy_compiler
        stats = {"sources": len(sources), "fresh_sources": 0, "new_structures": 0, "translate_s": 0.0, "nvrtc_s": 0.0,
                 "register_s": 0.0, "pool_workers": 0, "worker_translate_s_max": 0.0, "worker_nvrtc_s_max": 0.0}
        use_pool = (self.processes if processes is None else processes) > 1 and compile_pool.n_workers() > 0
        t0 = time.perf_counter()
        fresh = [s for s in dict.fromkeys(sources) if policy_compiler.cached(s) is None]
        stats["fresh_sources"] = len(fresh)
        self._errors = getattr(self, "_errors", {})
        if use_pool and len(fresh) >= self.POOL_SOURCES:
            nw = compile_pool.n_workers()
            per = -(-len(fresh) // nw)
            slices = [fresh[i:i + per] for i in range(0, len(fresh), per)]
            results, secs = compile_pool.run_jobs("translate", slices)
            stats["pool_workers"] = compile_pool.n_workers()
            stats["worker_translate_s_max"] = max(secs) if secs else 0.0
            for sl, res in zip(slices, results):
                if res is None:
                    continue  # that worker failed: these sources are translated in-process below
                for src, r in zip(sl, res):
                    if r[0] == "error":
                        self._errors[src] = r[1]
                    else:
                        policy_compiler.adopt(src, *r)
        out: List[Optional[CompiledPolicy]] = []
        for src in sources:
            if src in self._errors:
                out.append(None)
                continue
            try:
                out.append(compile_policy(src))
            except PolicyCompileError as e:
                self._errors[src] = str(e)
                out.append(None)
        stats["translate_s"] = time.perf_counter() - t0

        t0 = time.perf_counter()
        new = {}
        for cp in out:
            if cp is not None and cp.handle is None and cp.symbol not in new:
                new[cp.symbol] = (cp.cuda_source, cp.symbol, cp.flags)
        stats["new_structures"] = len(new)
        if new:
            structs = list(new.values())
            groups = [structs[i:i + self.COMPILE_BATCH] for i in range(0, len(structs), self.COMPILE_BATCH)]
            if use_pool and len(structs) >= self.POOL_STRUCTURES and len(groups) > 1:
                nw = max(1, compile_pool.n_workers())
                if len(groups) < nw:   # fewer programs than workers: smaller programs, all workers busy
                    g = max(4, -(-len(structs) // nw))
                    groups = [structs[i:i + g] for i in range(0, len(structs), g)]
                jobs = [groups[i::nw] for i in range(nw) if groups[i::nw]]
                _res, secs = compile_pool.run_jobs("nvrtc", jobs)
                stats["pool_workers"] = compile_pool.n_workers()
                stats["worker_nvrtc_s_max"] = max(secs) if secs else 0.0
                stats["nvrtc_s"] = time.perf_counter() - t0
                t0 = time.perf_counter()
                self._register_groups(groups, 1)   # cubins are in the shared cache: these are hits
            else:
                self._register_groups(groups, threads)
                stats["nvrtc_s"] = time.perf_counter() - t0
                t0 = time.perf_counter()
        stats["register_s"] = time.perf_counter() - t0
        self.last_compile_stats = stats
        return out

    def _register_groups(self, groups, threads: int) -> None:
        """ks_policy_compile_many per group (one NVRTC program each; a disk-cache hit when a worker compiled it)."""
        from . import compile_pool
        from .policy_compiler import _STRUCT_HANDLES
        import concurrent.futures as cf

        def one(group):
            return compile_pool.compile_group(self.lib, group)
        nt = threads if threads > 0 else min(len(groups), max(1, self.processes))
        if nt > 1 and len(groups) > 1:
            with cf.ThreadPoolExecutor(nt) as ex:   # ctypes releases the GIL inside the call
                done = list(ex.map(one, groups))
        else:
            done = [one(g) for g in groups]
        for group, (status, handles) in zip(groups, done):
            for (cuda_source, symbol, _flags), st, h in zip(group, status, handles):
                if st != 0:
                    raise nat.KsError(st, "NVRTC rejected translated policy:\n" +
                                      self.lib.ks_compile_log().decode(errors="replace") + "\n" + cuda_source)
                _STRUCT_HANDLES[symbol] = h

    def compile_error(self, source: str) -> Optional[str]:
        """Why `source` was declined by the translator (None if it was not)."""
        return getattr(self, "_errors", {}).get(source)

    def bind(self, policies: Sequence[CompiledPolicy], pin: bool = False) -> BoundPolicies:
        """Link (or find) a program that holds the structures of `policies` and lay out their float literals."""
        handles = []
        for p in policies:
            if p.handle is None:
                raise ValueError("policy was not compiled by BatchEvaluator.compile")
            handles.append(int(p.handle))
        need = frozenset(handles)
        hit = None
        for k in range(len(self._programs) - 1, -1, -1):
            if need <= self._programs[k][0]:
                hit = self._programs.pop(k)
                break
        if hit is None:
            order = sorted(need)
            if len(order) > nat.KS_MAX_TABLE:
                raise ValueError(f"{len(order)} distinct policy structures in one batch; a program holds {nat.KS_MAX_TABLE}")
            arr = (C.c_longlong * len(order))(*order)
            out = C.c_longlong(-1)
            nat.check(self.lib.ks_program_link(len(order), arr, C.byref(out)))
            hit = (need, {h: i for i, h in enumerate(order)}, out.value)
        self._programs.append(hit)   # most recently used last
        if pin:
            self._pinned.add(hit[2])
        while len(self._programs) > self.MAX_PROGRAMS:
            victims = [k for k, pr in enumerate(self._programs[:-1]) if pr[2] not in self._pinned]
            if not victims:
                break
            old = self._programs.pop(victims[0])
            self.lib.ks_program_release(old[2])
        slot_of = hit[1]
        slot = np.array([slot_of[h] for h in handles], dtype=np.int32)
        lens = np.array([len(p.consts) for p in policies], dtype=np.int64)
        off = np.concatenate([[0], np.cumsum(lens)[:-1]]) if len(policies) else np.zeros(0, np.int64)
        if lens.sum() >= 2 ** 31:
            raise ValueError("too many float literals in one batch")
        consts = np.array([c for p in policies for c in p.consts] or [0.0], dtype=np.float64)
        return BoundPolicies(hit[2], slot, off.astype(np.int32), consts)

    def _program(self, handles: Tuple[int, ...]) -> int:
        """Program holding exactly these structure handles, slot i = handles[i] (test hooks)."""
        arr = (C.c_longlong * len(handles))(*handles)
        out = C.c_longlong(-1)
        nat.check(self.lib.ks_program_link(len(handles), arr, C.byref(out)))
        self._pinned.add(out.value)
        return out.value

    # ------------------------------------------------------------------ evaluation
    def plan(self, n_episodes: int, want_placements: bool, max_pods: Optional[int] = None,
             teams_per_block: int = 0, warps_per_team: int = 0,
             episodes_per_policy: float = 1.0) -> nat.KsPlan:
        """Launch geometry (ks_plan).  The two-CTAs-per-SM kernel is chosen for batches of several waves whose table
        holds runs of same-policy episodes (policy-major order): warps of one SM then share a policy's code in the
        instruction cache and the extra resident warps hide latency; with a different policy per warp they thrash it
        (measured: +23 % on 512 policies x 16 traces, -6 % on 4,096 policies x 1 trace)."""
        p = nat.KsPlan()
        p.warps_per_team = warps_per_team or self.warps_per_team
        p.ctas_per_sm = self.ctas_per_sm
        if p.ctas_per_sm == 0 and p.warps_per_team <= 1 and self.cluster.n_nodes <= 64 and teams_per_block == 0 \
                and episodes_per_policy >= 8 and n_episodes >= 2 * 16 * self.sm_count:
            p.ctas_per_sm = 2
        p.n_episodes = n_episodes
        p.n_nodes = self.cluster.n_nodes
        p.max_pods = max_pods or self.max_pods
        p.want_placements = 1 if want_placements else 0
        p.teams_per_block = teams_per_block
        nat.check(self.lib.ks_plan(C.byref(p)))
        return p

    def run_episodes(self, policies: Sequence[CompiledPolicy], ep_policy: Sequence[int], ep_trace: Sequence[int],
                     want_placements: bool = False, max_events: int = 0, teams_per_block: int = 0,
                     stream=None) -> List[EpisodeOutcome]:
        """Evaluate episodes e = (policies[ep_policy[e]], traces[ep_trace[e]]) in one kernel launch."""
        torch = self._torch
        n_ep = len(ep_policy)
        if n_ep == 0:
            return []
        ep_policy = np.ascontiguousarray(ep_policy, dtype=np.int64)
        ep_trace = np.ascontiguousarray(ep_trace, dtype=np.int32)
        if ep_policy.min() < 0 or ep_policy.max() >= len(policies) or ep_trace.min() < 0 or ep_trace.max() >= len(self.traces):
            raise ValueError("episode table refers to an unknown policy or trace")
        bound = self.bind(policies)
        ep_slot = np.ascontiguousarray(bound.slot[ep_policy])
        ep_const = np.ascontiguousarray(bound.const_off[ep_policy])
        order = table_order(ep_slot, ep_const)
        if order is not None:
            ep_slot, ep_const, ep_trace = (np.ascontiguousarray(a[order]) for a in (ep_slot, ep_const, ep_trace))
        pods_per_ep = np.array([self.traces[t].n_pods for t in ep_trace], dtype=np.int64)
        plan = self.plan(n_ep, want_placements, int(pods_per_ep.max()), teams_per_block,
                         episodes_per_policy=n_ep / len(np.unique(ep_slot)))
        self.last_plan = plan
        st = stream or torch.cuda.current_stream(self.device)
        # copies, launch and read-back are all ordered on `st`
        with torch.cuda.device(self.device), torch.cuda.stream(st):
            d_pol = torch.from_numpy(ep_slot).to(self.device, non_blocking=True)
            d_cof = torch.from_numpy(ep_const).to(self.device, non_blocking=True)
            d_kf = torch.from_numpy(bound.consts).to(self.device, non_blocking=True)
            d_tr = torch.from_numpy(ep_trace).to(self.device, non_blocking=True)
            d_res = torch.empty(n_ep * nat.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=self.device)
            # one scratch buffer per stream: calls on different streams may overlap on the GPU
            key = int(st.cuda_stream)
            scratch = self._scratch.get(key)
            if scratch is None or scratch.numel() < plan.scratch_bytes:
                scratch = self._scratch[key] = torch.empty(int(plan.scratch_bytes), dtype=torch.uint8, device=self.device)
            d_place = d_off = None
            if want_placements:
                padded = (pods_per_ep + 1) // 2 * 2
                off = np.concatenate([[0], np.cumsum(padded)[:-1]]).astype(np.int64)
                d_off = torch.from_numpy(off).to(self.device, non_blocking=True)
                d_place = torch.empty(int(padded.sum()), dtype=torch.int64, device=self.device)
            b = nat.KsBatch()
            b.traces = self._trace_table.data_ptr()
            b.ep_policy = d_pol.data_ptr()
            b.ep_const = d_cof.data_ptr()
            b.consts = d_kf.data_ptr()
            b.ep_trace = d_tr.data_ptr()
            b.nconst = self._nconst.data_ptr()
            b.results = d_res.data_ptr()
            b.placements = d_place.data_ptr() if want_placements else None
            b.place_off = d_off.data_ptr() if want_placements else None
            b.scratch = scratch.data_ptr()
            for k in range(4):
                b.totals[k] = int(self._totals[k])
            b.max_events = max_events
            b.snapshot_interval = self.snapshot_interval
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record(st)
            nat.check(self.lib.ks_eval_batch(bound.program, C.byref(plan), C.byref(b), C.c_void_p(st.cuda_stream)))
            e1.record(st)
            h_res = d_res.to("cpu", non_blocking=True)
            h_place = d_place.to("cpu", non_blocking=True) if want_placements else None
            st.synchronize()
            res = np.frombuffer(h_res.numpy().tobytes(), dtype=nat.RESULT_DTYPE)
            self.last_kernel_ms = e0.elapsed_time(e1)
            place = h_place.numpy().view(np.uint64) if want_placements else None
        out = []
        for e in range(n_ep):
            r = res[e]
            results = None
            if r["num_snapshots"] > 0:
                results = EvaluationResults(float(r["avg_cpu_utilization"]), float(r["avg_memory_utilization"]),
                                            float(r["avg_gpu_count_utilization"]), float(r["avg_gpu_memory_utilization"]),
                                            float(r["gpu_fragmentation_score"]), int(r["num_snapshots"]),
                                            int(r["num_fragmentation_events"]))
            oc = EpisodeOutcome(int(r["status"]), float(r["policy_score"]), results, int(r["events"]), int(r["attempts"]),
                                int(r["scorer_calls"]), int(r["placed"]), int(r["max_nodes"]), int(r["max_waiting"]),
                                int(r["last_event_time"]), int(r["device_cycles"]))
            if want_placements:
                tr = self.traces[ep_trace[e]]
                rank = tr.packed()[0]
                rec = place[off[e]: off[e] + tr.n_pods][rank]   # rank-indexed -> pod_list order
                node = ((rec >> np.uint64(8)) & np.uint64(0xFFFF)).astype(np.int64)
                oc.final_creation_time = (rec >> np.uint64(24)).astype(np.int64)
                oc.node_index = np.where(node >= nat.KS_NODE_WAITING, -1, node)
                oc.waiting = node == nat.KS_NODE_WAITING
                oc.gpu_mask = (rec & np.uint64(0xFF)).astype(np.int64)
            out.append(oc)
        if order is not None:   # back to the caller's episode order
            unsorted: List[Optional[EpisodeOutcome]] = [None] * n_ep
            for k, e in enumerate(order):
                unsorted[int(e)] = out[k]
            out = unsorted
        return out

    # ------------------------------------------------------------------ test hook: raw scores of (pod, node) cases
    def debug_scores(self, policies: Sequence[CompiledPolicy], slot: int, pods: np.ndarray, ns: np.ndarray, nc: np.ndarray) -> np.ndarray:
        """Wrapped scores of policies[slot] for independent cases: pods [n, 6] int64, ns [n, KS_NS_FIELDS] int32,
        nc [n, KS_NC_FIELDS] int32 (ks_debug_score; field numbering of include/ks_abi.h).  -1 = the policy raised, -2 = left the supported subset."""
        torch = self._torch
        n = pods.shape[0]
        bound = self.bind(policies)
        d_p = torch.from_numpy(np.ascontiguousarray(pods, dtype=np.int64)).to(self.device)
        d_ns = torch.from_numpy(np.ascontiguousarray(ns, dtype=np.int32)).to(self.device)
        d_nc = torch.from_numpy(np.ascontiguousarray(nc, dtype=np.int32)).to(self.device)
        d_kf = torch.from_numpy(bound.consts).to(self.device)
        d_out = torch.empty(n, dtype=torch.int64, device=self.device)
        st = torch.cuda.current_stream(self.device)
        kf_ptr = d_kf.data_ptr() + 8 * int(bound.const_off[slot])
        nat.check(self.lib.ks_debug_score(bound.program, int(bound.slot[slot]), n, C.c_void_p(d_p.data_ptr()), C.c_void_p(d_ns.data_ptr()),
                                          C.c_void_p(d_nc.data_ptr()), C.c_void_p(kf_ptr), C.c_void_p(d_out.data_ptr()),
                                          C.c_void_p(st.cuda_stream)))
        return d_out.cpu().numpy()

    # ------------------------------------------------------------------ resident form (inputs already in HBM)
    def prepare_resident(self, policies: Sequence[CompiledPolicy], ep_policy: Sequence[int], ep_trace: Sequence[int],
                         teams_per_block: int = 0) -> "ResidentBatch":
        return ResidentBatch(self, policies, ep_policy, ep_trace, teams_per_block)

    # ------------------------------------------------------------------ host-buffer (end-to-end) form
    def prepare_host_batch(self, policies: Sequence[CompiledPolicy], ep_policy: Sequence[int], ep_trace: Sequence[int],
                           teams_per_block: int = 0) -> "HostBatch":
        """Stage a batch in pinned host memory for ks_eval_batch_host (everything the step needs crosses
        PCIe inside the call: traces, cluster table, episode table in; results out)."""
        return HostBatch(self, policies, ep_policy, ep_trace, teams_per_block)

    # ------------------------------------------------------------------ reference-shaped entry points
    def evaluate_population(self, codes: Sequence[str], trace_ids: Optional[Sequence[int]] = None,
                            on_unsupported: str = "zero") -> List[float]:
        """Fitness of every policy source of a generation: the batched replacement of the
        ProcessPoolExecutor block (funsearch_integration.py:535-562).  With several traces the fitness
        is the mean policy score over `trace_ids` (default: all resident traces).

        A policy that raises, runs short of GPUs or never places a pod scores 0, exactly like the reference's worker
        (:63-64, evaluator.py:108-110).  A policy OUTSIDE THE TRANSLATED SUBSET (the translator declined it, or it
        left int64 / real numbers at run time: status 'unsupported') is a different case: the reference would have
        executed it, so its true fitness is unknown here.  `on_unsupported` says what the list holds for those:
        "zero" (default; a warning counts them), "nan", or "raise".  `self.last_report` (PopulationReport) always
        lists them with the reason, so a caller can route them elsewhere."""
        import warnings
        if on_unsupported not in ("zero", "nan", "raise"):
            raise ValueError("on_unsupported must be 'zero', 'nan' or 'raise'")
        trace_ids = list(range(len(self.traces))) if trace_ids is None else list(trace_ids)
        compiled = self.compile(codes)
        uniq, slot = [], {}
        for cp in compiled:
            if cp is not None and id(cp) not in slot:
                slot[id(cp)] = len(uniq)
                uniq.append(cp)
        ep_pol, ep_tr, owner = [], [], []
        for i, cp in enumerate(compiled):
            if cp is None:
                continue
            for t in trace_ids:
                ep_pol.append(slot[id(cp)])
                ep_tr.append(t)
                owner.append(i)
        scores = [0.0] * len(codes)
        status = [nat.ST_OK if cp is not None else nat.ST_UNSUPPORTED for cp in compiled]
        messages = {i: "translator: " + (self.compile_error(codes[i]) or "declined") for i, cp in enumerate(compiled) if cp is None}
        if ep_pol:
            outs = self.run_episodes(uniq, ep_pol, ep_tr)
            acc = [0.0] * len(codes)
            for o, i in zip(outs, owner):
                acc[i] += o.policy_score if o.status == nat.ST_OK else 0.0
                if o.status != nat.ST_OK and status[i] == nat.ST_OK:
                    status[i] = o.status
                    messages[i] = "episode status: " + o.status_name
            # any failed episode of a policy is an exception in the reference's worker: the policy scores 0
            scores = [a / len(trace_ids) if status[i] == nat.ST_OK else 0.0 for i, a in enumerate(acc)]
        not_eval = [i for i, st in enumerate(status) if st == nat.ST_UNSUPPORTED]
        self.last_report = PopulationReport(list(scores), status, not_eval, messages,
                                            len({cp.symbol for cp in compiled if cp is not None}))
        if not_eval:
            if on_unsupported == "raise":
                raise nat.KsError(nat.ST_UNSUPPORTED, f"{len(not_eval)} of {len(codes)} policies are outside the translated subset: "
                                  + "; ".join(f"#{i}: {messages[i]}" for i in not_eval[:5]))
            if on_unsupported == "nan":
                for i in not_eval:
                    scores[i] = float("nan")
            else:
                warnings.warn(f"{len(not_eval)} of {len(codes)} policies are outside the translated subset and were NOT "
                              "evaluated (score 0 here, unlike the reference, which would run them); see last_report",
                              RuntimeWarning, stacklevel=2)
        return scores

    def evaluate_policy_standalone(self, policy_data: Tuple[int, str]) -> Tuple[int, str, float]:
        """Same signature and error behaviour as funsearch_integration.py:30-64."""
        idx, code = policy_data
        try:
            return (idx, code, self.evaluate_population([code], [0])[0])
        except Exception:  # noqa: BLE001 -- the reference swallows everything into score 0
            return (idx, code, 0)

    def evaluate_policy_full(self, code: str) -> Optional[float]:
        """SimpleFunSearch._evaluate_policy_full (funsearch_integration.py:433-459): the score, or None (after printing
        the reason) where the reference's in-process variant catches an exception."""
        try:
            score = self.evaluate_population([code], [0])[0]
            rep = self.last_report
            if rep.status[0] != nat.ST_OK:
                print(f"Error evaluating policy: {rep.messages.get(0, 'failed')}")
                return None
            return score
        except Exception as e:  # noqa: BLE001
            print(f"Error evaluating policy: {e}")
            return None


def _episode_tables(ev: BatchEvaluator, policies, ep_policy, ep_trace, pin=True):
    """Bind + structure-major episode table shared by the resident and the host-buffer batch."""
    ep_policy = np.ascontiguousarray(ep_policy, dtype=np.int64)
    ep_trace = np.ascontiguousarray(ep_trace, dtype=np.int32)
    if len(ep_policy) == 0 or ep_policy.min() < 0 or ep_policy.max() >= len(policies) or ep_trace.min() < 0 \
            or ep_trace.max() >= len(ev.traces):
        raise ValueError("episode table is empty or refers to an unknown policy or trace")
    bound = ev.bind(policies, pin=pin)
    ep_slot = np.ascontiguousarray(bound.slot[ep_policy])
    ep_const = np.ascontiguousarray(bound.const_off[ep_policy])
    order = table_order(ep_slot, ep_const)
    if order is not None:
        ep_slot, ep_const, ep_trace = (np.ascontiguousarray(a[order]) for a in (ep_slot, ep_const, ep_trace))
    return bound, ep_slot, ep_const, ep_trace, order


class ResidentBatch:
    """A batch whose episode table, result buffer and scratch already live in HBM: launch() is exactly one
    ks_eval_batch call (one memset + one kernel), nothing else."""

    def __init__(self, ev: BatchEvaluator, policies, ep_policy, ep_trace, teams_per_block=0):
        torch = ev._torch
        self.ev = ev
        self.n_ep = len(ep_policy)
        bound, ep_slot, ep_const, ep_trace, self.order = _episode_tables(ev, policies, ep_policy, ep_trace)
        self.prog = bound.program
        self.n_structures = int(len(np.unique(ep_slot)))
        if self.order is not None:   # the device table is structure-major; results() undoes it
            self.inverse = np.argsort(self.order)
            self._d_inverse = torch.from_numpy(self.inverse).to(ev.device)
        pods_per_ep = [ev.traces[t].n_pods for t in ep_trace]
        self.plan = ev.plan(self.n_ep, False, max(pods_per_ep), teams_per_block,
                            episodes_per_policy=self.n_ep / self.n_structures)
        self.d_pol = torch.from_numpy(ep_slot).to(ev.device)
        self.d_cof = torch.from_numpy(ep_const).to(ev.device)
        self.d_kf = torch.from_numpy(bound.consts).to(ev.device)
        self.d_tr = torch.from_numpy(ep_trace).to(ev.device)
        self.d_res = torch.zeros(self.n_ep * nat.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=ev.device)
        self.d_scratch = torch.empty(int(self.plan.scratch_bytes), dtype=torch.uint8, device=ev.device)
        b = nat.KsBatch()
        b.traces = ev._trace_table.data_ptr()
        b.ep_policy = self.d_pol.data_ptr()
        b.ep_const = self.d_cof.data_ptr()
        b.consts = self.d_kf.data_ptr()
        b.ep_trace = self.d_tr.data_ptr()
        b.nconst = ev._nconst.data_ptr()
        b.results = self.d_res.data_ptr()
        b.placements = None
        b.place_off = None
        b.scratch = self.d_scratch.data_ptr()
        for k in range(4):
            b.totals[k] = int(ev._totals[k])
        b.max_events = 0
        b.snapshot_interval = ev.snapshot_interval
        self.batch = b

    def launch(self, stream=None) -> None:
        torch = self.ev._torch
        st = stream or torch.cuda.current_stream(self.ev.device)
        nat.check(self.ev.lib.ks_eval_batch(self.prog, C.byref(self.plan), C.byref(self.batch), C.c_void_p(st.cuda_stream)))

    def results(self) -> np.ndarray:
        res = np.frombuffer(self.d_res.cpu().numpy().tobytes(), dtype=nat.RESULT_DTYPE)
        return res if self.order is None else res[self.inverse]

    def scores_tensor(self):
        """Per-episode policy_score as a device fp64 tensor (a strided view of the result records)."""
        torch = self.ev._torch
        rec = self.d_res.view(torch.float64).view(self.n_ep, nat.RESULT_DTYPE.itemsize // 8)
        col = rec[:, nat.RESULT_DTYPE.fields["policy_score"][1] // 8]
        return col if self.order is None else col[self._d_inverse]


class HostBatch:
    """One batch held in pinned host memory + a device workspace; run() = one ks_eval_batch_host call."""

    def __init__(self, ev: BatchEvaluator, policies, ep_policy, ep_trace, teams_per_block=0):
        torch = ev._torch
        self.ev = ev
        self.n_ep = len(ep_policy)
        bound, ep_slot, ep_const, ep_trace, self.order = _episode_tables(ev, policies, ep_policy, ep_trace)
        self.prog = bound.program
        if self.order is not None:   # the table that crosses PCIe is structure-major; run() undoes it
            self.inverse = np.argsort(self.order)
        pods_per_ep = [ev.traces[t].n_pods for t in ep_trace]
        self.plan = ev.plan(self.n_ep, False, max(pods_per_ep), teams_per_block,
                            episodes_per_policy=self.n_ep / len(np.unique(ep_slot)))

        def pinned(a: np.ndarray):
            t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
            return t
        self._keep = []
        nconst, totals = ev.cluster.packed()
        self.h_nconst = pinned(nconst)
        self.h_pol = pinned(ep_slot)
        self.h_cof = pinned(ep_const)
        self.h_kf = pinned(bound.consts)
        self.h_tr = pinned(ep_trace)
        self.h_res = torch.empty(self.n_ep * nat.RESULT_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
        self.h_traces = (nat.KsHostTrace * len(ev.traces))()
        for i, tr in enumerate(ev.traces):
            _r, podrec, heap0, place0 = tr.packed()
            bufs = (pinned(podrec), pinned(heap0.view(np.int64)), pinned(place0.view(np.int64)))
            self._keep.append(bufs)
            self.h_traces[i].podrec = bufs[0].data_ptr()
            self.h_traces[i].heap0 = bufs[1].data_ptr()
            self.h_traces[i].place0 = bufs[2].data_ptr()
            self.h_traces[i].n_pods = tr.n_pods
        hb = nat.KsHostBatch()
        hb.traces = self.h_traces
        hb.n_traces = len(ev.traces)
        hb.ep_policy = self.h_pol.data_ptr()
        hb.ep_const = self.h_cof.data_ptr()
        hb.consts = self.h_kf.data_ptr()
        hb.n_consts = int(bound.consts.size)
        hb.ep_trace = self.h_tr.data_ptr()
        hb.nconst = self.h_nconst.data_ptr()
        hb.results = self.h_res.data_ptr()
        hb.placements = None
        hb.place_off = None
        hb.placements_len = 0
        for k in range(4):
            hb.totals[k] = int(totals[k])
        hb.max_events = 0
        hb.snapshot_interval = ev.snapshot_interval
        self.hb = hb
        need = C.c_longlong(0)
        nat.check(ev.lib.ks_host_workspace_bytes(C.byref(self.plan), C.byref(hb), C.byref(need)))
        self.workspace = torch.empty(int(need.value), dtype=torch.uint8, device=ev.device)
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def run(self, stream=None) -> np.ndarray:
        """Host buffers in, host results out; returns the KsResult records (numpy structured array)."""
        torch = self.ev._torch
        st = stream or torch.cuda.current_stream(self.ev.device)
        up, down = C.c_longlong(0), C.c_longlong(0)
        nat.check(self.ev.lib.ks_eval_batch_host(self.prog, C.byref(self.plan), C.byref(self.hb),
                                                 C.c_void_p(self.workspace.data_ptr()), self.workspace.numel(),
                                                 C.c_void_p(st.cuda_stream), C.byref(up), C.byref(down)))
        self.h2d_bytes, self.d2h_bytes = up.value, down.value
        res = np.frombuffer(self.h_res.numpy(), dtype=nat.RESULT_DTYPE)
        return res if self.order is None else res[self.inverse]


def evaluate_population(codes: Sequence[str], cluster: Optional[ClusterSpec] = None,
                        traces: Optional[Sequence[Trace]] = None, device: int = 0) -> List[float]:
    """One-call form on the reference's default workload (parse_workload() defaults,
    benchmarks/parser.py:117-122: gpu_models_filtered.csv x openb_pod_list_default.csv)."""
    cluster = cluster or ClusterSpec.load("gpu16")
    traces = traces or [Trace.load("default")]
    return BatchEvaluator(cluster, traces, device).evaluate_population(codes)
