"""Host-side workloads: pod traces and clusters in the reference's vocabulary, packed for the GPU.

`Trace` and `ClusterSpec` accept either plain arrays, the bundled .npz files (converted from the
reference's OpenB CSVs by tests/golden/make_golden.py) or *the reference's own objects* --
anything with the field names of simulator/entities.py:4-43 (duck-typed), so
`Trace.from_pods(parser.parse_pods())` and `ClusterSpec.from_cluster(parser.parse_cluster())`
work on a reference checkout without conversion code.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Iterable, Optional, Sequence

import numpy as np

from . import _native as nat

DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def id_ranks(pod_ids: Sequence[str]) -> np.ndarray:
    """rank of each pod_id in Python string order (Event.__lt__, simulator/event_simulator.py:16-17).

    Ranks come from sorting the strings, never from a numeric suffix: 'openb-pod-10000' sorts
    before 'openb-pod-9999'."""
    ids = list(pod_ids)
    order = sorted(range(len(ids)), key=ids.__getitem__)
    for a, b in zip(order, order[1:]):
        if ids[a] == ids[b]:
            raise ValueError(f"duplicate pod_id {ids[a]!r}: the event order of the reference is undefined for equal ids")
    rank = np.empty(len(ids), dtype=np.int32)
    rank[np.asarray(order, dtype=np.int64)] = np.arange(len(ids), dtype=np.int32)
    return rank


@dataclass
class Trace:
    """A pod list in pod_list order (simulator/entities.py:30-43 fields)."""
    pod_ids: list
    cpu_milli: np.ndarray
    memory_mib: np.ndarray
    num_gpu: np.ndarray
    gpu_milli: np.ndarray
    creation_time: np.ndarray
    duration_time: np.ndarray
    name: str = ""

    def __post_init__(self):
        self.cpu_milli = np.ascontiguousarray(self.cpu_milli, dtype=np.int32)
        self.memory_mib = np.ascontiguousarray(self.memory_mib, dtype=np.int32)
        self.num_gpu = np.ascontiguousarray(self.num_gpu, dtype=np.int32)
        self.gpu_milli = np.ascontiguousarray(self.gpu_milli, dtype=np.int32)
        self.creation_time = np.ascontiguousarray(self.creation_time, dtype=np.int64)
        self.duration_time = np.ascontiguousarray(self.duration_time, dtype=np.int64)
        self.pod_ids = [str(s) for s in self.pod_ids]
        n = len(self.pod_ids)
        for f in ("cpu_milli", "memory_mib", "num_gpu", "gpu_milli", "creation_time", "duration_time"):
            if getattr(self, f).shape != (n,):
                raise ValueError(f"Trace.{f} must have one entry per pod")
        self._packed = None

    @property
    def n_pods(self) -> int:
        return len(self.pod_ids)

    @classmethod
    def from_pods(cls, pods: Iterable, name: str = "") -> "Trace":
        pods = list(pods)
        return cls([p.pod_id for p in pods], [p.cpu_milli for p in pods], [p.memory_mib for p in pods],
                   [p.num_gpu for p in pods], [p.gpu_milli for p in pods], [p.creation_time for p in pods],
                   [p.duration_time for p in pods], name)

    @classmethod
    def load(cls, name_or_path: str) -> "Trace":
        path = name_or_path if os.path.exists(name_or_path) else os.path.join(DATA_DIR, f"trace_{name_or_path}.npz")
        z = np.load(path)
        return cls(list(z["pod_ids"]), z["cpu_milli"], z["memory_mib"], z["num_gpu"], z["gpu_milli"],
                   z["creation_time"], z["duration_time"], os.path.basename(path))

    def head(self, n: int) -> "Trace":
        return Trace(self.pod_ids[:n], self.cpu_milli[:n], self.memory_mib[:n], self.num_gpu[:n], self.gpu_milli[:n],
                     self.creation_time[:n], self.duration_time[:n], f"{self.name}[:{n}]")

    def packed(self):
        """(id_rank, podrec, heap0, place0) host arrays in the device layout (ks_trace_pack)."""
        if self._packed is None:
            n = self.n_pods
            rank = id_ranks(self.pod_ids)
            podrec = np.zeros(max(n, 1) * nat.KS_POD_INTS, dtype=np.int32)
            # 16-byte vector copies on the device read pairs of entries: keep one spare slot
            heap0 = np.zeros(n + 2, dtype=np.uint64)
            place0 = np.zeros(n + 2, dtype=np.uint64)
            nat.check(nat.lib().ks_trace_pack(n, nat.ptr(self.cpu_milli), nat.ptr(self.memory_mib), nat.ptr(self.num_gpu),
                                              nat.ptr(self.gpu_milli), nat.ptr(self.creation_time),
                                              nat.ptr(self.duration_time), nat.ptr(rank), nat.ptr(podrec),
                                              nat.ptr(heap0), nat.ptr(place0)))
            self._packed = (rank, podrec, heap0, place0)
        return self._packed


@dataclass
class ClusterSpec:
    """Nodes in nodes_dict insertion order (simulator/entities.py:4-26 fields)."""
    node_ids: list
    cpu_total: np.ndarray
    mem_total: np.ndarray
    num_gpus: np.ndarray
    gpu_milli_total: np.ndarray  # [n, 8]
    gpu_mem_total: np.ndarray    # [n, 8]
    name: str = ""

    def __post_init__(self):
        self.node_ids = [str(s) for s in self.node_ids]
        n = len(self.node_ids)
        self.cpu_total = np.ascontiguousarray(self.cpu_total, dtype=np.int32)
        self.mem_total = np.ascontiguousarray(self.mem_total, dtype=np.int32)
        self.num_gpus = np.ascontiguousarray(self.num_gpus, dtype=np.int32)
        self.gpu_milli_total = np.ascontiguousarray(self.gpu_milli_total, dtype=np.int32).reshape(n, 8)
        self.gpu_mem_total = np.ascontiguousarray(self.gpu_mem_total, dtype=np.int32).reshape(n, 8)
        self._packed = None

    @property
    def n_nodes(self) -> int:
        return len(self.node_ids)

    @property
    def n_pad(self) -> int:
        return (self.n_nodes + 31) // 32 * 32

    @classmethod
    def from_cluster(cls, cluster, name: str = "") -> "ClusterSpec":
        nodes = list(cluster.nodes_dict.values())
        n = len(nodes)
        gmil = np.zeros((n, 8), dtype=np.int32)
        gmem = np.zeros((n, 8), dtype=np.int32)
        for i, nd in enumerate(nodes):
            if len(nd.gpus) > 8:
                raise ValueError(f"node {nd.node_id} has {len(nd.gpus)} GPUs; at most 8 are supported")
            # The replay starts every episode from a pristine copy of the cluster (what the reference's callers pass:
            # a deepcopy of the parsed cluster, funsearch_integration.py:46-48).  A cluster that is already partly
            # used, or whose gpu_left disagrees with its GPU list, would be replayed differently by the reference.
            if getattr(nd, "gpu_left", len(nd.gpus)) != len(nd.gpus) or nd.cpu_milli_left != nd.cpu_milli_total \
                    or nd.memory_mib_left != nd.memory_mib_total or any(g.gpu_milli_left != g.gpu_milli_total for g in nd.gpus):
                raise ValueError(f"node {nd.node_id} is not pristine (its *_left fields differ from its totals, or gpu_left "
                                 "from len(gpus)): pass the cluster as parsed, before any replay")
            for j, g in enumerate(nd.gpus):
                gmil[i, j] = g.gpu_milli_total
                gmem[i, j] = g.memory_mib_total
        return cls([nd.node_id for nd in nodes], [nd.cpu_milli_total for nd in nodes],
                   [nd.memory_mib_total for nd in nodes], [len(nd.gpus) for nd in nodes], gmil, gmem, name)

    @classmethod
    def load(cls, name_or_path: str) -> "ClusterSpec":
        path = name_or_path if os.path.exists(name_or_path) else os.path.join(DATA_DIR, f"cluster_{name_or_path}.npz")
        z = np.load(path)
        return cls(list(z["node_ids"]), z["cpu_total"], z["mem_total"], z["num_gpus"], z["gpu_milli_total"],
                   z["gpu_mem_total"], os.path.basename(path))

    def head(self, n: int) -> "ClusterSpec":
        return ClusterSpec(self.node_ids[:n], self.cpu_total[:n], self.mem_total[:n], self.num_gpus[:n],
                           self.gpu_milli_total[:n], self.gpu_mem_total[:n], f"{self.name}[:{n}]")

    def packed(self):
        """(nconst, totals) host arrays in the device layout (ks_cluster_pack)."""
        if self._packed is None:
            nconst = np.zeros(nat.KS_NC_FIELDS * self.n_pad, dtype=np.int32)
            totals = np.zeros(4, dtype=np.int64)
            nat.check(nat.lib().ks_cluster_pack(self.n_nodes, nat.ptr(self.cpu_total), nat.ptr(self.mem_total),
                                                nat.ptr(self.num_gpus), nat.ptr(self.gpu_milli_total),
                                                nat.ptr(self.gpu_mem_total), nat.ptr(nconst), nat.ptr(totals)))
            self._packed = (nconst, totals)
        return self._packed


def available_traces() -> list:
    return sorted(f[len("trace_"):-4] for f in os.listdir(DATA_DIR) if f.startswith("trace_") and f.endswith(".npz"))


def bootstrap_trace(trace: Trace, seed: int) -> Trace:
    """A seeded variant of `trace` (BASELINE config 4's "seeds": the simulator itself has no randomness, SURVEY.md
    section 0): the arrival times and pod ids stay, the resource rows (cpu, mem, num_gpu, gpu_milli, duration) are
    resampled with replacement from the trace's own rows.  seed 0 returns the trace unchanged."""
    if seed == 0:
        return trace
    rng = np.random.Generator(np.random.PCG64(seed))
    rows = rng.integers(0, trace.n_pods, size=trace.n_pods)
    return Trace(trace.pod_ids, trace.cpu_milli[rows], trace.memory_mib[rows], trace.num_gpu[rows], trace.gpu_milli[rows],
                 trace.creation_time, trace.duration_time[rows], f"{trace.name}#seed{seed}")


def synthetic_trace(n_pods: int, seed: int = 0, base: Optional[Trace] = None, load_factor: float = 1.0,
                    name: Optional[str] = None) -> Trace:
    """Alibaba-shaped synthetic trace: rows (cpu, mem, num_gpu, gpu_milli,
    duration) are bootstrapped from the empirical joint distribution of `base` (default: the bundled
    default trace), inter-arrival gaps are resampled from its empirical gaps and divided by
    `load_factor`.  Ids are fixed-width so string order equals numeric order."""
    base = base or Trace.load("default")
    rng = np.random.Generator(np.random.PCG64(seed))
    rows = rng.integers(0, base.n_pods, size=n_pods)
    order = np.argsort(base.creation_time, kind="stable")
    gaps = np.diff(base.creation_time[order])
    gaps = gaps[gaps >= 0]
    g = gaps[rng.integers(0, len(gaps), size=n_pods)].astype(np.float64) / float(load_factor)
    ctime = np.floor(np.cumsum(g)).astype(np.int64)
    width = max(7, len(str(max(n_pods - 1, 1))))
    ids = [f"synth-pod-{i:0{width}d}" for i in range(n_pods)]
    return Trace(ids, base.cpu_milli[rows], base.memory_mib[rows], base.num_gpu[rows], base.gpu_milli[rows], ctime,
                 base.duration_time[rows], name or f"synthetic{n_pods}_s{seed}_lf{load_factor:g}")
