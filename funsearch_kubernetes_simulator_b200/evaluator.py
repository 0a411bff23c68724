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
    invariant_check: int = 0                  # debug build, status 6: (failing check << 32) | node index

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


def table_order(*keys: np.ndarray) -> Optional[np.ndarray]:
    """Stable permutation that sorts the episode table by `keys` (most significant first: module, structure slot,
    policy), or None when it already is sorted.  The kernel hands each CTA a contiguous range of the table, so episodes
    that run the same device function sit side by side on one SM and share its code in the instruction cache
    (csrc/ks_sim_kernel.cu, work distribution)."""
    n = len(keys[0])
    if n < 2:
        return None
    order = np.lexsort(tuple(np.asarray(k) for k in reversed(keys)))   # lexsort: last key is the primary one
    if np.array_equal(order, np.arange(n)):
        return None
    return order


@dataclass
class Module:
    """One linked and loaded program: the replay kernel + a set of policy structures."""
    program: int
    slot_of: dict            # structure handle -> slot in the module's dispatch table
    last_use: int = 0


@dataclass
class BoundPolicies:
    """A list of compiled policies bound to loaded modules: per policy the module, the table slot (structure) and the
    start of its float literals; the literals of all of them back to back."""
    modules: List[Module]
    part: np.ndarray        # int32 [n_policies]: index into `modules`
    slot: np.ndarray        # int32 [n_policies]
    const_off: np.ndarray   # int32 [n_policies]
    consts: np.ndarray      # float64

    @property
    def program(self) -> int:
        if len(self.modules) != 1:
            raise ValueError("policies are spread over several modules")
        return self.modules[0].program


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

    # structures per NVRTC program (ks_set_compile_batch): a worker's share of a generation (35-45 structures) is one program,
    # so the headers are parsed once per worker (16 -> 48: compile 1.39-1.57 -> 1.24-1.37 s per fresh generation)
    COMPILE_BATCH = int(os.environ.get("KS_COMPILE_BATCH", "48"))
    POOL_SOURCES = 48           # fresh sources from which translating in the pool pays off
    POOL_STRUCTURES = 24        # new structures from which NVRTC + link in the pool pays off
    MAX_MODULES = 128           # loaded modules kept (least recently used ones are released)
    MAX_PARTS = 32              # modules one batch may be spread over; beyond, the smallest contributions are merged
    CONSOLIDATE_PARTS = 1       # a batch spread over more modules than this is relinked into ONE module in the background

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
        self._modules: List[Module] = []
        self._module_of: dict = {}     # structure handle -> newest module holding it
        self._pinned: set = set()
        self._clock = 0
        self._keep: set = set()        # modules the batch being bound lives in: never evicted while it is bound
        self._bg_thread = None         # background relink of the last generation's structures into one module
        self._bg_done: list = []       # finished relinks: (handles in table order, program), adopted by the next bind()
        self._streams: list = []
        self._errors: dict = {}
        self.last_plan = None
        self.last_kernel_ms = None
        self.last_compile_stats: dict = {}
        self.last_report: Optional[PopulationReport] = None
        self.sm_count = torch.cuda.get_device_properties(self.device).multi_processor_count
        self.warps_per_team = 0   # 0 = planner's choice; set to force a team width
        self.ctas_per_sm = 0      # 0 = planner's choice; 1 = latency kernel, 2 = dense (throughput) kernel
        self.module_structures = 0  # test knob: at most this many structures per linked module (0 = no limit)
        self.pack_by_cost = True    # single-wave batches: cut the table into per-SM ranges by predicted cost
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
    def compile(self, sources: Sequence[str], threads: int = 0, processes: Optional[int] = None) -> List[Optional[CompiledPolicy]]:
        """Python policy sources -> compiled policies (structure function + float literals).
        A source outside the supported subset yields None (its episodes are reported as not evaluated).
        A generation of fresh sources is translated by the pool of worker processes started with the evaluator
        (`processes=0` keeps everything in this process); only structures never seen before reach NVRTC, in groups
        of COMPILE_BATCH, again on the pool, and each worker also links what it compiled with the replay kernel:
        this process only loads the finished modules."""
        import time
        from . import compile_pool, policy_compiler
        stats = {"sources": len(sources), "fresh_sources": 0, "new_structures": 0, "translate_s": 0.0, "nvrtc_link_s": 0.0,
                 "register_load_s": 0.0, "pool_workers": 0, "worker_translate_s_max": 0.0, "worker_nvrtc_link_s_max": 0.0,
                 "modules_loaded": 0}
        use_pool = (self.processes if processes is None else processes) > 1 and compile_pool.n_workers() > 0
        t0 = time.perf_counter()
        fresh = [s for s in dict.fromkeys(sources) if policy_compiler.cached(s) is None and s not in self._errors]
        stats["fresh_sources"] = len(fresh)
        if use_pool and len(fresh) >= self.POOL_SOURCES:
            nw = max(1, min(compile_pool.n_workers(), self.processes if processes is None else processes))
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
                # Deal the structures to the workers longest-first by source size (NVRTC time grows with it), then cut
                # each worker's share into programs of at most COMPILE_BATCH: every worker ends at about the same time.
                nw = max(1, min(compile_pool.n_workers(), self.processes if processes is None else processes,
                                -(-len(structs) // 4)))
                shares: List[list] = [[] for _ in range(nw)]
                load = [0] * nw
                for st_ in sorted(structs, key=lambda x: -len(x[0])):
                    k = load.index(min(load))
                    shares[k].append(st_)
                    load[k] += len(st_[0]) + 2000
                per_worker = []
                for k, sh in enumerate(shares):
                    # big and small structures alternate inside a module: table neighbours share an SM, and eight
                    # 6 KB policies side by side thrash its instruction cache (measured: slowest episode +30 %)
                    sh = [sh[(i // 2) if i % 2 == 0 else len(sh) - 1 - i // 2] for i in range(len(sh))]
                    shares[k] = sh
                    n_prog = -(-len(sh) // self.COMPILE_BATCH)
                    size = -(-len(sh) // n_prog)
                    per_worker.append([sh[i:i + size] for i in range(0, len(sh), size)])
                groups = [g for gs in per_worker for g in gs]
                images = [os.path.join(compile_pool.cache_dir(), f"module_{os.getpid()}_{self._clock}_{k}.cubin")
                          for k in range(len(per_worker))]
                self._clock += 1
                res, secs = compile_pool.run_jobs("nvrtc_link", [(gs, img) for gs, img in zip(per_worker, images)])
                stats["pool_workers"] = compile_pool.n_workers()
                stats["worker_nvrtc_link_s_max"] = max(secs) if secs else 0.0
                stats["worker_nvrtc_link_s_mean"] = sum(secs) / len(secs) if secs else 0.0
                stats["nvrtc_link_s"] = time.perf_counter() - t0
                t0 = time.perf_counter()
                self._register_groups(groups, 1)   # cubins are in the shared cache: these are hits
                for gs, img, r in zip(per_worker, images, res):
                    if r is None or not r.get("linked") or not os.path.exists(img):
                        continue   # bind() links whatever has no module yet
                    handles = [policy_compiler._STRUCT_HANDLES[sym] for g in gs for (_src, sym, _fl) in g]
                    with open(img, "rb") as f:
                        blob = f.read()
                    os.unlink(img)
                    arr = (C.c_longlong * len(handles))(*handles)
                    prog = C.c_longlong(-1)
                    nat.check(self.lib.ks_program_load_image(blob, len(blob), len(handles), arr, C.byref(prog)))
                    self._add_module(Module(prog.value, {h: i for i, h in enumerate(handles)}))
                    stats["modules_loaded"] += 1
            else:
                self._register_groups(groups, threads)
                stats["nvrtc_link_s"] = time.perf_counter() - t0
                t0 = time.perf_counter()
        stats["register_load_s"] = time.perf_counter() - t0
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
        return self._errors.get(source)

    # ------------------------------------------------------------------ modules
    def _add_module(self, m: Module) -> Module:
        self._clock += 1
        m.last_use = self._clock
        self._modules.append(m)
        for h in m.slot_of:
            self._module_of[h] = m
        while len(self._modules) > self.MAX_MODULES:
            victims = [x for x in self._modules if x.program not in self._pinned and x is not m and id(x) not in self._keep]
            if not victims:
                break
            old = min(victims, key=lambda x: x.last_use)
            self._modules.remove(old)
            for h in list(old.slot_of):
                if self._module_of.get(h) is old:
                    del self._module_of[h]
                    for other in self._modules:      # another module may still hold the structure
                        if h in other.slot_of:
                            self._module_of[h] = other
                            break
            self._torch.cuda.synchronize(self.device)   # no kernel of the module may still be running
            self.lib.ks_program_release(old.program)
        return m

    def _adopt_consolidated(self) -> None:
        """Modules linked in the background become the home of their structures (the small per-worker modules they
        replace age out of the cache)."""
        while self._bg_done:
            order, prog = self._bg_done.pop()
            self._add_module(Module(prog, {h: i for i, h in enumerate(order)}))
            # modules all of whose structures now live elsewhere will never be bound again: release them
            for old in [m for m in self._modules if m.program not in self._pinned and id(m) not in self._keep
                        and all(self._module_of.get(h) is not m for h in m.slot_of)]:
                self._modules.remove(old)
                self._torch.cuda.synchronize(self.device)
                self.lib.ks_program_release(old.program)

    def _consolidate_async(self, handles: Sequence[int]) -> None:
        """Relink `handles` into one module on a helper thread (nvJitLink and the module load release the GIL).  The
        compile workers hand a fresh generation over as one module per worker, which is what makes the first
        evaluation fast; between generations -- while the caller's LLM writes the next one -- the generation is merged,
        so that a later batch finds last generation's survivors in one module, not in sixteen."""
        import threading
        if self._bg_thread is not None and self._bg_thread.is_alive():
            return
        order = sorted(set(handles))
        if len(order) > nat.KS_MAX_TABLE or len(order) < 2:
            return

        def work():
            arr = (C.c_longlong * len(order))(*order)
            out = C.c_longlong(-1)
            if self.lib.ks_program_link(len(order), arr, C.byref(out)) == 0:
                self._bg_done.append((order, out.value))
        self._bg_thread = threading.Thread(target=work, daemon=True)
        self._bg_thread.start()
        if not getattr(self, "_atexit", False):   # the helper thread must not be inside the linker when the process exits
            import atexit
            import weakref
            ref = weakref.ref(self)
            atexit.register(lambda: ref() is not None and ref()._bg_thread is not None and ref()._bg_thread.join())
            self._atexit = True

    def wait_background(self) -> None:
        """Block until a background relink (if any) has finished and adopt its module."""
        if self._bg_thread is not None:
            self._bg_thread.join()
        self._adopt_consolidated()

    def _link_module(self, handles: Sequence[int]) -> Module:
        order = sorted(set(handles))
        arr = (C.c_longlong * len(order))(*order)
        out = C.c_longlong(-1)
        nat.check(self.lib.ks_program_link(len(order), arr, C.byref(out)))
        return self._add_module(Module(out.value, {h: i for i, h in enumerate(order)}))

    def bind(self, policies: Sequence[CompiledPolicy], pin: bool = False, single_module: bool = False) -> BoundPolicies:
        """Find (or link) modules that hold the structures of `policies` and lay out their float literals.  Structures
        keep the module they were first linked into, so a generation that mostly re-tunes constants, or repeats last
        generation's elites, links only what is new; a batch spread over more than MAX_PARTS modules (or asked to be
        in one) has its structures relinked into a single module."""
        handles = []
        for p in policies:
            if p.handle is None:
                raise ValueError("policy was not compiled by BatchEvaluator.compile")
            handles.append(int(p.handle))
        need = sorted(set(handles))
        self._adopt_consolidated()
        cap = min(nat.KS_MAX_TABLE, self.module_structures or nat.KS_MAX_TABLE)
        for _attempt in range(4):
            # the modules this batch already lives in are the most recently used ones from here on: linking below may
            # evict old modules, but not these
            self._clock += 1
            self._keep = set()
            for h in need:
                m = self._module_of.get(h)
                if m is not None:
                    m.last_use = self._clock
                    self._keep.add(id(m))
            missing = [h for h in need if h not in self._module_of]
            if not missing:
                break
            for i in range(0, len(missing), cap):
                self._link_module(missing[i:i + cap])
        used: dict = {}
        for h in need:
            used.setdefault(id(self._module_of[h]), []).append(h)
        if single_module and len(used) > 1 and not self.module_structures:
            for i in range(0, len(need), nat.KS_MAX_TABLE):
                self._link_module(need[i:i + nat.KS_MAX_TABLE])   # newest module wins in _module_of
        elif len(used) > self.MAX_PARTS and not self.module_structures:
            # Elites and re-tuned survivors of many earlier generations: each old module contributes a few structures.
            # The smallest contributions are relinked into one module (cost ~ what is moved) until 3/4 of the limit.
            moved, left = [], len(used)
            for hs in sorted(used.values(), key=len):
                if left + 1 <= self.MAX_PARTS * 3 // 4:
                    break
                moved.extend(hs)
                left -= 1
            for i in range(0, len(moved), nat.KS_MAX_TABLE):
                self._link_module(moved[i:i + nat.KS_MAX_TABLE])
        mods: List[Module] = []
        index = {}
        self._clock += 1
        part = np.zeros(len(handles), dtype=np.int32)
        slot = np.zeros(len(handles), dtype=np.int32)
        for k, h in enumerate(handles):
            m = self._module_of[h]
            if id(m) not in index:
                index[id(m)] = len(mods)
                mods.append(m)
                m.last_use = self._clock
                if pin:
                    self._pinned.add(m.program)
            part[k] = index[id(m)]
            slot[k] = m.slot_of[h]
        self._keep = set()
        if len(mods) > self.CONSOLIDATE_PARTS and not self.module_structures:
            self._consolidate_async(need)
        lens = np.array([len(p.consts) for p in policies], dtype=np.int64)
        off = np.concatenate([[0], np.cumsum(lens)[:-1]]) if len(policies) else np.zeros(0, np.int64)
        if lens.sum() >= 2 ** 31:
            raise ValueError("too many float literals in one batch")
        consts = np.array([c for p in policies for c in p.consts] or [0.0], dtype=np.float64)
        return BoundPolicies(mods, part, slot, off.astype(np.int32), consts)

    def _program(self, handles: Tuple[int, ...]) -> int:
        """Program holding exactly these structure handles, slot i = handles[i] (test hooks)."""
        arr = (C.c_longlong * len(handles))(*handles)
        out = C.c_longlong(-1)
        nat.check(self.lib.ks_program_link(len(handles), arr, C.byref(out)))
        self._pinned.add(out.value)
        return out.value

    def _side_streams(self, n: int):
        torch = self._torch
        while len(self._streams) < n:
            self._streams.append(torch.cuda.Stream(device=self.device))
        return self._streams[:n]

    # ------------------------------------------------------------------ evaluation
    def plan(self, n_episodes: int, want_placements: bool, max_pods: Optional[int] = None,
             teams_per_block: int = 0, warps_per_team: int = 0,
             episodes_per_policy: float = 1.0, allow_dense: bool = True) -> nat.KsPlan:
        """Launch geometry (ks_plan).  The two-CTAs-per-SM kernel is chosen for batches of several waves whose table
        holds runs of same-structure episodes: warps of one SM then share a policy's code in the
        instruction cache and the extra resident warps hide latency; with a different function per warp they thrash it
        (measured: +23 % on 512 policies x 16 traces, -6 % on 4,096 policies x 1 trace)."""
        p = nat.KsPlan()
        p.warps_per_team = warps_per_team or self.warps_per_team
        p.ctas_per_sm = self.ctas_per_sm
        if p.ctas_per_sm == 0 and p.warps_per_team <= 1 and self.cluster.n_nodes <= 64 and teams_per_block == 0 \
                and allow_dense and episodes_per_policy >= 8 and n_episodes >= 2 * 16 * self.sm_count:
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
        """Evaluate episodes e = (policies[ep_policy[e]], traces[ep_trace[e]]): one kernel launch per module the
        policies' structures live in (one launch when they share a module), all in flight together."""
        if len(ep_policy) == 0:
            return []
        rb = ResidentBatch(self, policies, ep_policy, ep_trace, teams_per_block, want_placements=want_placements,
                           max_events=max_events, pin=False)
        self.last_plan = rb.plans[0]
        torch = self._torch
        st = stream or torch.cuda.current_stream(self.device)
        with torch.cuda.device(self.device):
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record(st)
            rb.launch(st)
            e1.record(st)
            res = rb.results(st)
            self.last_kernel_ms = e0.elapsed_time(e1)
            place = rb.placements(st) if want_placements else None
        out = []
        for e in range(rb.n_ep):
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
            if oc.status == nat.ST_INVARIANT:
                oc.invariant_check = int(r["reserved"][0])
            if want_placements:
                tr = self.traces[int(ep_trace[e])]
                rank = tr.packed()[0]
                rec = place[e][rank]   # rank-indexed -> pod_list order
                node = ((rec >> np.uint64(8)) & np.uint64(0xFFFF)).astype(np.int64)
                oc.final_creation_time = (rec >> np.uint64(24)).astype(np.int64)
                oc.node_index = np.where(node >= nat.KS_NODE_WAITING, -1, node)
                oc.waiting = node == nat.KS_NODE_WAITING
                oc.gpu_mask = (rec & np.uint64(0xFF)).astype(np.int64)
            out.append(oc)
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
        prog = bound.modules[int(bound.part[slot])].program
        nat.check(self.lib.ks_debug_score(prog, int(bound.slot[slot]), n, C.c_void_p(d_p.data_ptr()), C.c_void_p(d_ns.data_ptr()),
                                          C.c_void_p(d_nc.data_ptr()), C.c_void_p(kf_ptr), C.c_void_p(d_out.data_ptr()),
                                          C.c_void_p(st.cuda_stream)))
        return d_out.cpu().numpy()

    # ------------------------------------------------------------------ resident form (inputs already in HBM)
    def prepare_resident(self, policies: Sequence[CompiledPolicy], ep_policy: Sequence[int], ep_trace: Sequence[int],
                         teams_per_block: int = 0, single_module: Optional[bool] = None, max_events: int = 0,
                         costs=None) -> "ResidentBatch":
        """`costs` (optional, one per episode): predicted relative cost, e.g. device cycles or events of an earlier replay
        of the same or the parent policy; single-wave batches are packed onto the SMs by it (_pack_single_wave)."""
        return ResidentBatch(self, policies, ep_policy, ep_trace, teams_per_block, single_module=single_module,
                             max_events=max_events, costs=costs)

    # ------------------------------------------------------------------ host-buffer (end-to-end) form
    def prepare_host_batch(self, policies: Sequence[CompiledPolicy], ep_policy: Sequence[int], ep_trace: Sequence[int],
                           teams_per_block: int = 0, single_module: Optional[bool] = None, max_events: int = 0,
                           costs=None) -> "HostBatch":
        """Stage a batch in pinned host memory for ks_eval_batch_host (everything the step needs crosses
        PCIe inside the call: traces, cluster table, episode table in; results out)."""
        return HostBatch(self, policies, ep_policy, ep_trace, teams_per_block, single_module=single_module,
                         max_events=max_events, costs=costs)

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
            torch = self._torch
            rb = ResidentBatch(self, uniq, ep_pol, ep_tr, pin=False)
            st_ = torch.cuda.current_stream(self.device)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st_)
            rb.launch(st_)
            e1.record(st_)
            res = rb.results(st_)
            self.last_kernel_ms = e0.elapsed_time(e1)
            self.last_plan = rb.plans[0]
            self.last_results = res
            own = np.asarray(owner, dtype=np.int64)
            ok_ep = res["status"] == nat.ST_OK
            acc = np.bincount(own, weights=np.where(ok_ep, res["policy_score"], 0.0), minlength=len(codes))
            # any failed episode of a policy is an exception in the reference's worker: the policy scores 0
            worst = np.zeros(len(codes), dtype=np.int64)
            np.maximum.at(worst, own, np.where(ok_ep, 0, res["status"]).astype(np.int64))
            for i in np.nonzero(worst)[0]:
                # a real failure (exception, GPU shortfall) in any episode makes 0 the reference's answer too; only a
                # policy whose failures are all 'unsupported' counts as not evaluated
                first = int(res["status"][(own == i) & ~ok_ep].min())
                status[int(i)] = first
                messages[int(i)] = "episode status: " + nat.STATUS_NAMES.get(first, str(first))
            scores = [float(acc[i]) / len(trace_ids) if status[i] == nat.ST_OK else 0.0 for i in range(len(codes))]
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



class _Tables:
    """Bind + the episode table sorted module-major, then structure-major, then policy-major; cut into one part per
    module.  Shared by the resident and the host-buffer batch."""

    def __init__(self, ev: BatchEvaluator, policies, ep_policy, ep_trace, teams_per_block, want_placements, pin,
                 single_module, costs=None):
        ep_policy = np.ascontiguousarray(ep_policy, dtype=np.int64)
        ep_trace = np.ascontiguousarray(ep_trace, dtype=np.int32)
        self.n_ep = len(ep_policy)
        if self.n_ep == 0 or ep_policy.min() < 0 or ep_policy.max() >= len(policies) or ep_trace.min() < 0 \
                or ep_trace.max() >= len(ev.traces):
            raise ValueError("episode table is empty or refers to an unknown policy or trace")
        if single_module is None:   # batches of several waves: one module, one tail, and the two-CTAs-per-SM kernel
            single_module = self.n_ep >= 2 * 16 * ev.sm_count   # (the relink is small next to such a run)
        self.bound = bound = ev.bind(policies, pin=pin, single_module=single_module)
        part = bound.part[ep_policy]
        self.ep_slot = bound.slot[ep_policy]
        self.ep_const = bound.const_off[ep_policy]
        self.ep_trace = ep_trace
        self.order = table_order(part, self.ep_slot, self.ep_const)
        if self.order is not None:
            part = part[self.order]
            self.ep_slot, self.ep_const, self.ep_trace = (np.ascontiguousarray(a[self.order])
                                                          for a in (self.ep_slot, self.ep_const, self.ep_trace))
            self.inverse = np.argsort(self.order)
        else:
            self.ep_slot, self.ep_const = np.ascontiguousarray(self.ep_slot), np.ascontiguousarray(self.ep_const)
        # predicted cost of each episode (sorted order): the caller's (e.g. cycles or events of an earlier replay, of the
        # parent policy), else a proxy: size of the policy's code x pods of the trace
        if costs is not None:
            cost = np.asarray(costs, dtype=np.float64)
            if cost.shape != (self.n_ep,):
                raise ValueError("one cost per episode")
        else:
            size = np.array([len(p.cuda_source) for p in policies], dtype=np.float64)
            cost = size[ep_policy] * np.array([ev.traces[t].n_pods for t in ep_trace], dtype=np.float64)
        self.cost = cost if self.order is None else cost[self.order]
        self.range_begin = None
        self.n_structures = int(len(np.unique(part.astype(np.int64) * nat.KS_MAX_TABLE + self.ep_slot)))
        self.pods_per_ep = np.array([ev.traces[t].n_pods for t in self.ep_trace], dtype=np.int64)
        # one part per module: [start, end) of the sorted table
        cuts = [0] + [int(i) for i in np.nonzero(np.diff(part))[0] + 1] + [self.n_ep]
        self.parts = [(bound.modules[int(part[a])], a, b) for a, b in zip(cuts[:-1], cuts[1:])]
        # Geometry.  One module: ks_plan's own choice.  Several modules: the SMs are dealt to the parts in proportion to
        # their episodes (at least one CTA each), with as many teams per CTA as a single launch of the whole table would
        # use, so that all CTAs are resident together and every SM carries about the same number of episodes.  (A plan
        # may carry any block count: the kernel cuts the part's table into one balanced range per CTA.)
        n_parts = len(self.parts)
        self.plans = []
        if n_parts == 1:
            p = ev.plan(self.n_ep, want_placements, int(self.pods_per_ep.max()), teams_per_block,
                        episodes_per_policy=self.n_ep / self.n_structures)
            self.plans.append(p)
            if p.warps_per_team == 1 and p.ctas_per_sm == 1 and teams_per_block == 0 and ev.sm_count < self.n_ep <= 8 * ev.sm_count \
                    and not want_placements and ev.pack_by_cost:
                self._pack_single_wave(ev, want_placements)
            return
        whole = ev.plan(self.n_ep, False, 1, teams_per_block, allow_dense=False)
        tpb = whole.teams_per_block
        sizes = [b - a for _m, a, b in self.parts]
        sms = ev.sm_count
        if -(-self.n_ep // tpb) <= sms:      # everything fits in one wave: spread it over all SMs
            blocks = [max(1, (n * sms) // self.n_ep) for n in sizes]
            while sum(blocks) > sms and max(blocks) > 1:
                k = max(range(n_parts), key=lambda i: blocks[i] / sizes[i] if blocks[i] > 1 else 0.0)
                blocks[k] -= 1
            while sum(blocks) < sms:
                k = max(range(n_parts), key=lambda i: sizes[i] / blocks[i])
                if sizes[k] / blocks[k] <= 1.0:
                    break
                blocks[k] += 1
            blocks = [max(b, -(-n // tpb)) for b, n in zip(blocks, sizes)]   # never more than tpb episodes per CTA at once
        else:
            blocks = [min(sms, -(-n // tpb)) for n in sizes]
        for (_m, a, b), nb in zip(self.parts, blocks):
            p = ev.plan(b - a, want_placements, int(self.pods_per_ep[a:b].max()), tpb,
                        episodes_per_policy=self.n_ep / self.n_structures, allow_dense=False)
            p.blocks = int(nb)
            p.scratch_bytes = nat.KS_SCRATCH_HEADER + p.scratch_per_team * p.blocks * p.teams_per_block
            self.plans.append(p)


def _pack_single_wave(self, ev, want_placements):
    """A generation that fits the GPU in one wave is as long as its slowest episode, and an episode is slower the more
    CODE its SM has to keep in the instruction cache (profiles/README.md).  The table is therefore cut into one explicit
    range per SM by LPT on the predicted cost: episodes in order of decreasing cost, each to the SM with the smallest
    load so far, at most 8 per SM; the kernel replays exactly these ranges (no stealing).  Measured on configs[1]
    (tools/pack_probe.py): with the default cost -- the size of the policy's code, which needs no earlier measurement --
    116.7 -> 106.9 ms; cycles or events measured by an earlier replay are worse predictors (111 / 117 ms), and giving
    the heaviest episodes SMs of their own is worse than the default cut (146 ms).  Result order is restored by the
    callers through `order` / `inverse` as before."""
    import heapq
    n, sms = self.n_ep, ev.sm_count
    by = np.argsort(-self.cost, kind="stable")
    # LPT with at most 8 per SM: the heaviest episode left goes to the SM with the smallest predicted load so far
    heap = [(0.0, k) for k in range(sms)]
    bins = [[] for _ in range(sms)]
    for e in by:
        while True:
            load, k = heapq.heappop(heap)
            if len(bins[k]) < 8:
                break
        bins[k].append(int(e))
        if len(bins[k]) < 8:
            heapq.heappush(heap, (load + float(self.cost[e]), k))
    bins = [bn for bn in bins if bn]
    sizes = [len(bn) for bn in bins]
    by_cost = np.array([e for bn in bins for e in bn], dtype=np.int64)
    c = self.cost[by_cost]
    hi = max(sum(self.cost[e] for e in bn) for bn in bins) / max(float(self.cost.max()), 1e-30)
    plan = ev.plan(n, want_placements, int(self.pods_per_ep.max()), 8, episodes_per_policy=1.0, allow_dense=False)
    plan.blocks = len(sizes)
    plan.scratch_bytes = nat.KS_SCRATCH_HEADER + plan.scratch_per_team * plan.blocks * plan.teams_per_block
    self.plans = [plan]
    self.range_begin = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    # the table in packed order
    self.ep_slot, self.ep_const, self.ep_trace = (np.ascontiguousarray(a[by_cost]) for a in (self.ep_slot, self.ep_const, self.ep_trace))
    self.pods_per_ep = self.pods_per_ep[by_cost]
    self.cost = self.cost[by_cost]
    self.order = by_cost if self.order is None else self.order[by_cost]
    self.inverse = np.argsort(self.order)
    self.predicted_makespan_over_heaviest = hi


_Tables._pack_single_wave = _pack_single_wave


class ResidentBatch:
    """A batch whose episode table, result buffer and scratch already live in HBM: launch() is one ks_eval_batch call
    (one memset + one kernel) per module the batch's structures live in, all in flight together, nothing else."""

    def __init__(self, ev: BatchEvaluator, policies, ep_policy, ep_trace, teams_per_block=0, want_placements=False,
                 max_events=0, pin=True, single_module=None, costs=None):
        torch = ev._torch
        self.ev = ev
        t = self.tables = _Tables(ev, policies, ep_policy, ep_trace, teams_per_block, want_placements, pin, single_module, costs)
        self.n_ep = t.n_ep
        self.order = t.order
        self.n_structures = t.n_structures
        self.plans = t.plans
        self.plan = t.plans[0]
        self.n_modules = len(t.parts)
        self.prog = t.parts[0][0].program
        if t.order is not None:   # the device table is sorted; results() undoes it
            self.inverse = t.inverse
            self._d_inverse = torch.from_numpy(t.inverse).to(ev.device)
        rec = nat.RESULT_DTYPE.itemsize
        with torch.cuda.device(ev.device):
            self.d_pol = torch.from_numpy(t.ep_slot).to(ev.device)
            self.d_cof = torch.from_numpy(t.ep_const).to(ev.device)
            self.d_kf = torch.from_numpy(t.bound.consts).to(ev.device)
            self.d_tr = torch.from_numpy(t.ep_trace).to(ev.device)
            self.d_res = torch.zeros(self.n_ep * rec, dtype=torch.uint8, device=ev.device)
            self.d_ranges = torch.from_numpy(t.range_begin).to(ev.device) if t.range_begin is not None else None
            self.d_place = self.d_off = None
            if want_placements:
                padded = (t.pods_per_ep + 1) // 2 * 2
                self.place_off = np.concatenate([[0], np.cumsum(padded)[:-1]]).astype(np.int64)
                self.d_off = torch.from_numpy(self.place_off).to(ev.device)
                self.d_place = torch.empty(int(padded.sum()), dtype=torch.int64, device=ev.device)
            self.d_scratch = [torch.empty(int(p.scratch_bytes), dtype=torch.uint8, device=ev.device) for p in self.plans]
        self.batches = []
        for (m, a, b), plan, scratch in zip(t.parts, self.plans, self.d_scratch):
            kb = nat.KsBatch()
            kb.traces = ev._trace_table.data_ptr()
            kb.ep_policy = self.d_pol.data_ptr() + 4 * a
            kb.ep_const = self.d_cof.data_ptr() + 4 * a
            kb.consts = self.d_kf.data_ptr()
            kb.ep_trace = self.d_tr.data_ptr() + 4 * a
            kb.range_begin = self.d_ranges.data_ptr() if self.d_ranges is not None else None
            kb.nconst = ev._nconst.data_ptr()
            kb.results = self.d_res.data_ptr() + rec * a
            kb.placements = self.d_place.data_ptr() if want_placements else None
            kb.place_off = self.d_off.data_ptr() + 8 * a if want_placements else None
            kb.scratch = scratch.data_ptr()
            for k in range(4):
                kb.totals[k] = int(ev._totals[k])
            kb.max_events = max_events
            kb.snapshot_interval = ev.snapshot_interval
            self.batches.append((m.program, plan, kb))
        self.batch = self.batches[0][2]
        self._fork = torch.cuda.Event()
        self._joins = [torch.cuda.Event() for _ in self.batches]

    def launch(self, stream=None) -> None:
        """Enqueue the batch on `stream` (default: torch's current stream).  With several modules the launches go to
        side streams that fork from and join back into `stream`, so they run concurrently and `stream` order holds."""
        torch = self.ev._torch
        st = stream or torch.cuda.current_stream(self.ev.device)
        lib = self.ev.lib
        if len(self.batches) == 1:
            prog, plan, kb = self.batches[0]
            nat.check(lib.ks_eval_batch(prog, C.byref(plan), C.byref(kb), C.c_void_p(st.cuda_stream)))
            return
        side = self.ev._side_streams(len(self.batches))
        self._fork.record(st)
        for (prog, plan, kb), s, j in zip(self.batches, side, self._joins):
            s.wait_event(self._fork)
            nat.check(lib.ks_eval_batch(prog, C.byref(plan), C.byref(kb), C.c_void_p(s.cuda_stream)))
            j.record(s)
        for j in self._joins:
            st.wait_event(j)

    def _to_host(self, dev_tensor, stream):
        torch = self.ev._torch
        st = stream or torch.cuda.current_stream(self.ev.device)
        with torch.cuda.stream(st):
            h = dev_tensor.to("cpu", non_blocking=True)
        st.synchronize()
        return h

    def results(self, stream=None) -> np.ndarray:
        """KsResult records in the caller's episode order (waits for `stream`)."""
        res = np.frombuffer(self._to_host(self.d_res, stream).numpy().tobytes(), dtype=nat.RESULT_DTYPE)
        return res if self.order is None else res[self.inverse]

    def placements(self, stream=None) -> list:
        """Per episode (caller's order): the rank-indexed placement records."""
        place = self._to_host(self.d_place, stream).numpy().view(np.uint64)
        t = self.tables
        rows = [place[self.place_off[k]: self.place_off[k] + t.pods_per_ep[k]] for k in range(self.n_ep)]
        return rows if self.order is None else [rows[int(self.inverse[e])] for e in range(self.n_ep)]

    def status_tensor(self):
        """Per-episode status word as a device int32 tensor (caller's order)."""
        torch = self.ev._torch
        col = self.d_res.view(torch.int32).view(self.n_ep, nat.RESULT_DTYPE.itemsize // 4)[:, 0]
        return col if self.order is None else col[self._d_inverse]

    def scores_tensor(self):
        """Per-episode policy_score as a device fp64 tensor (a strided view of the result records)."""
        torch = self.ev._torch
        rec = self.d_res.view(torch.float64).view(self.n_ep, nat.RESULT_DTYPE.itemsize // 8)
        col = rec[:, nat.RESULT_DTYPE.fields["policy_score"][1] // 8]
        return col if self.order is None else col[self._d_inverse]


class HostBatch:
    """One batch held in pinned host memory + device workspaces; run() = one ks_eval_batch_host call per module the
    batch's structures live in (concurrently, each on its own stream)."""

    def __init__(self, ev: BatchEvaluator, policies, ep_policy, ep_trace, teams_per_block=0, single_module=None, max_events=0,
                 costs=None):
        torch = ev._torch
        self.ev = ev
        t = self.tables = _Tables(ev, policies, ep_policy, ep_trace, teams_per_block, False, True, single_module, costs)
        self.n_ep = t.n_ep
        self.order = t.order
        if t.order is not None:   # the table that crosses PCIe is sorted; run() undoes it
            self.inverse = t.inverse
        self.plans = t.plans
        self.plan = t.plans[0]
        rec = nat.RESULT_DTYPE.itemsize

        def pinned(a: np.ndarray):
            return torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        self._keep = []
        nconst, totals = ev.cluster.packed()
        self.h_nconst = pinned(nconst)
        self.h_pol = pinned(t.ep_slot)
        self.h_cof = pinned(t.ep_const)
        self.h_kf = pinned(t.bound.consts)
        self.h_tr = pinned(t.ep_trace)
        self.h_ranges = pinned(t.range_begin) if t.range_begin is not None else None
        self.h_res = torch.empty(self.n_ep * rec, dtype=torch.uint8).pin_memory()
        self.h_traces = (nat.KsHostTrace * len(ev.traces))()
        for i, tr in enumerate(ev.traces):
            _r, podrec, heap0, place0 = tr.packed()
            bufs = (pinned(podrec), pinned(heap0.view(np.int64)), pinned(place0.view(np.int64)))
            self._keep.append(bufs)
            self.h_traces[i].podrec = bufs[0].data_ptr()
            self.h_traces[i].heap0 = bufs[1].data_ptr()
            self.h_traces[i].place0 = bufs[2].data_ptr()
            self.h_traces[i].n_pods = tr.n_pods
        self.calls = []
        for (m, a, b), plan in zip(t.parts, self.plans):
            hb = nat.KsHostBatch()
            hb.traces = self.h_traces
            hb.n_traces = len(ev.traces)
            hb.ep_policy = self.h_pol.data_ptr() + 4 * a
            hb.ep_const = self.h_cof.data_ptr() + 4 * a
            hb.consts = self.h_kf.data_ptr()
            hb.n_consts = int(t.bound.consts.size)
            hb.ep_trace = self.h_tr.data_ptr() + 4 * a
            hb.range_begin = self.h_ranges.data_ptr() if self.h_ranges is not None else None
            hb.nconst = self.h_nconst.data_ptr()
            hb.results = self.h_res.data_ptr() + rec * a
            hb.placements = None
            hb.place_off = None
            hb.placements_len = 0
            for k in range(4):
                hb.totals[k] = int(totals[k])
            hb.max_events = max_events
            hb.snapshot_interval = ev.snapshot_interval
            need = C.c_longlong(0)
            nat.check(ev.lib.ks_host_workspace_bytes(C.byref(plan), C.byref(hb), C.byref(need)))
            ws = torch.empty(int(need.value), dtype=torch.uint8, device=ev.device)
            self.calls.append((m.program, plan, hb, ws))
        self.hb = self.calls[0][2]
        self.workspace = self.calls[0][3]
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._pool = None

    def _one(self, call, cuda_stream):
        prog, plan, hb, ws = call
        up, down = C.c_longlong(0), C.c_longlong(0)
        rc = self.ev.lib.ks_eval_batch_host(prog, C.byref(plan), C.byref(hb), C.c_void_p(ws.data_ptr()), ws.numel(),
                                            C.c_void_p(cuda_stream), C.byref(up), C.byref(down))
        return rc, (self.ev.lib.ks_last_error().decode(errors="replace") if rc else ""), up.value, down.value

    def run(self, stream=None) -> np.ndarray:
        """Host buffers in, host results out; returns the KsResult records (numpy structured array)."""
        torch = self.ev._torch
        st = stream or torch.cuda.current_stream(self.ev.device)
        if len(self.calls) == 1:
            outs = [self._one(self.calls[0], st.cuda_stream)]
        else:
            # each call copies, launches and waits for its own stream: run them from threads (ctypes releases the GIL)
            import concurrent.futures as cf
            if self._pool is None:
                self._pool = cf.ThreadPoolExecutor(len(self.calls))
            st.synchronize()
            side = self.ev._side_streams(len(self.calls))
            outs = list(self._pool.map(lambda cs: self._one(cs[0], cs[1].cuda_stream), zip(self.calls, side)))
        for rc, msg, _u, _d in outs:
            if rc:
                raise nat.KsError(rc, msg)
        self.h2d_bytes, self.d2h_bytes = sum(o[2] for o in outs), sum(o[3] for o in outs)
        res = np.frombuffer(self.h_res.numpy(), dtype=nat.RESULT_DTYPE)
        return res if self.order is None else res[self.inverse]


def evaluate_population(codes: Sequence[str], cluster: Optional[ClusterSpec] = None,
                        traces: Optional[Sequence[Trace]] = None, device: int = 0) -> List[float]:
    """One-call form on the reference's default workload (parse_workload() defaults,
    benchmarks/parser.py:117-122: gpu_models_filtered.csv x openb_pod_list_default.csv)."""
    cluster = cluster or ClusterSpec.load("gpu16")
    traces = traces or [Trace.load("default")]
    return BatchEvaluator(cluster, traces, device).evaluate_population(codes)
