"""Object-level drop-ins around the batched evaluator.

* `GpuKubernetesSimulator` keeps the reference's object protocol for ONE episode:
  KubernetesSimulator(cluster, pod_list, ...).run_schedule() mutating Pod.assigned_node /
  assigned_gpus / creation_time in place (simulator/main.py:28-72, 126-138) and exposing max_nodes and
  get_evaluation_results().
* validate_invariants=True (simulator/main.py:201-272) has two counterparts: with the debug build of the library
  (build.py --invariants, selected by KS_B200_LIB) the KERNEL runs the reference's checks on its live state after every
  mutation and ends the episode on the first violation; with the production build `check_invariants` replays the
  placement log on the host and checks that no node or GPU was ever over-committed and that the accounting returns to
  the totals.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np

from .evaluator import BatchEvaluator, EpisodeOutcome, EvaluationResults
from .workload import ClusterSpec, Trace


def write_back(outcome: EpisodeOutcome, pods: List, cluster: ClusterSpec) -> None:
    """Fill the reference's Pod objects from a placement log: assigned_node ('' if never placed),
    assigned_gpus (ascending GPU indices; the reference stores them in best-fit order, the set is the same)
    and the retry-mutated creation_time."""
    if outcome.node_index is None:
        raise ValueError("run the episode with want_placements=True")
    for i, pod in enumerate(pods):
        ni = int(outcome.node_index[i])
        pod.creation_time = int(outcome.final_creation_time[i])
        if ni >= 0:
            pod.assigned_node = cluster.node_ids[ni]
            m = int(outcome.gpu_mask[i])
            pod.assigned_gpus = [g for g in range(8) if (m >> g) & 1]
        else:
            pod.assigned_node = ""
            pod.assigned_gpus = []


INVARIANT_MESSAGES = {
    1: "has negative CPU remaining", 2: "has negative memory remaining", 3: "has negative GPU count remaining",
    4: "CPU remaining exceeds total", 5: "memory remaining exceeds total", 6: "GPU remaining exceeds total",
    7: "has a GPU with negative milli remaining", 8: "has a GPU whose milli remaining exceeds total",
    9: "CPU accounting error", 10: "memory accounting error", 11: "GPU accounting error", 12: "GPU milli accounting error"}


def write_back_cluster(outcome: EpisodeOutcome, cluster_obj) -> None:
    """Node / GPU `*_left` after a completed replay (simulator/main.py:83-96, 126-128): every placed pod has also been
    deleted by the time the event heap is empty, so every resource is back at its total."""
    for node in cluster_obj.nodes_dict.values():
        node.cpu_milli_left = node.cpu_milli_total
        node.memory_mib_left = node.memory_mib_total
        node.gpu_left = len(node.gpus)
        for g in node.gpus:
            g.gpu_milli_left = g.gpu_milli_total


def check_invariants(outcome: EpisodeOutcome, cluster: ClusterSpec, trace: Trace) -> None:
    """Raises ValueError if, at any time of the replay, a node's cpu/memory/GPU count or a GPU's milli went
    negative or above its total (simulator/main.py:201-272).  Pods occupy [final creation time, +duration)."""
    if outcome.node_index is None:
        raise ValueError("run the episode with want_placements=True")
    rank = trace.packed()[0]
    placed = np.nonzero(outcome.node_index >= 0)[0]
    t0 = outcome.final_creation_time[placed]
    t1 = t0 + trace.duration_time[placed]
    # event order of the simulator: (time, id_rank); a pod's deletion never precedes its own creation
    ev_t = np.concatenate([t0, t1])
    ev_r = np.concatenate([rank[placed], rank[placed]])
    ev_kind = np.concatenate([np.zeros(len(placed), np.int8), np.ones(len(placed), np.int8)])
    ev_pod = np.concatenate([placed, placed])
    order = np.lexsort((ev_kind, ev_r, ev_t))
    cpu = cluster.cpu_total.astype(np.int64).copy()
    mem = cluster.mem_total.astype(np.int64).copy()
    cnt = cluster.num_gpus.astype(np.int64).copy()
    gml = cluster.gpu_milli_total.astype(np.int64).copy()
    for k in order:
        p = int(ev_pod[k])
        n = int(outcome.node_index[p])
        sign = -1 if ev_kind[k] == 0 else 1
        cpu[n] += sign * int(trace.cpu_milli[p])
        mem[n] += sign * int(trace.memory_mib[p])
        cnt[n] += sign * int(trace.num_gpu[p])
        m = int(outcome.gpu_mask[p])
        for g in range(8):
            if (m >> g) & 1:
                gml[n, g] += sign * int(trace.gpu_milli[p])
                if not 0 <= gml[n, g] <= cluster.gpu_milli_total[n, g]:
                    raise ValueError(f"Node {cluster.node_ids[n]} GPU {g} milli out of range: {gml[n, g]}")
        if cpu[n] < 0 or mem[n] < 0 or cnt[n] < 0:
            raise ValueError(f"Node {cluster.node_ids[n]} over-committed at t={int(ev_t[k])}: cpu {cpu[n]} mem {mem[n]} gpus {cnt[n]}")
        if cpu[n] > cluster.cpu_total[n] or mem[n] > cluster.mem_total[n] or cnt[n] > cluster.num_gpus[n]:
            raise ValueError(f"Node {cluster.node_ids[n]} accounting exceeds totals at t={int(ev_t[k])}")
    if not (np.array_equal(cpu, cluster.cpu_total) and np.array_equal(mem, cluster.mem_total)
            and np.array_equal(cnt, cluster.num_gpus) and np.array_equal(gml, cluster.gpu_milli_total)):
        raise ValueError("resources do not return to the totals after all deletions")


class GpuKubernetesSimulator:
    """One-episode facade with the reference's object protocol.

        sim = GpuKubernetesSimulator(cluster_obj, pod_list, policy_source)
        sim.run_schedule()                # mutates the Pod objects like the reference does
        sim.get_evaluation_results(); sim.get_policy_score(); sim.max_nodes

    `cluster_obj` / `pod_list` are the reference's own entities (or anything with the same field names).
    Batches of episodes should use BatchEvaluator directly; this class exists for call sites that hold
    entity objects (tests/test_scheduler.py-style drivers, simulator/main.py users)."""

    def __init__(self, cluster, pod_list: List, policy_source: str, validate_invariants: bool = False,
                 device: int = 0, evaluator: Optional[BatchEvaluator] = None):
        self.cluster_obj = cluster
        self.pod_list = pod_list
        self.policy_source = policy_source
        self.validate_invariants = validate_invariants
        self.cluster = ClusterSpec.from_cluster(cluster)
        self.trace = Trace.from_pods(pod_list)
        self._ev = evaluator or BatchEvaluator(self.cluster, [self.trace], device)
        self.max_nodes = 0
        self.waiting_pods: List = []     # pods that failed at least once and were never placed (main.py:42, 116-117)
        self.outcome: Optional[EpisodeOutcome] = None

    def run_schedule(self) -> None:
        compiled = self._ev.compile([self.policy_source])
        if compiled[0] is None:
            raise ValueError("policy source is outside the supported subset")
        self.outcome = self._ev.run_episodes(compiled, [0], [0], want_placements=True)[0]
        if self.outcome.status == 1:
            raise RuntimeError("the policy raised during the replay (the reference's worker would score it 0)")
        if self.outcome.status == 2:
            raise ValueError("Not enough GPUs available on the chosen node")   # simulator/main.py:164-165
        if self.outcome.status == 6:   # debug build of the library: the kernel's own _validate_cluster_invariants failed
            code, node = self.outcome.invariant_check >> 32, self.outcome.invariant_check & 0xFFFFFFFF
            raise ValueError(f"Node {self.cluster.node_ids[node]} {INVARIANT_MESSAGES.get(code, 'violates an invariant')}")
        if self.outcome.status != 0:
            raise RuntimeError(f"episode ended with status {self.outcome.status_name}")
        write_back(self.outcome, self.pod_list, self.cluster)
        write_back_cluster(self.outcome, self.cluster_obj)
        self.waiting_pods = [p for p, w in zip(self.pod_list, self.outcome.waiting) if w]
        self.max_nodes = self.outcome.max_nodes
        if self.validate_invariants and not (self._ev.lib.ks_build_flags() & 2):
            check_invariants(self.outcome, self.cluster, self.trace)

    def get_evaluation_results(self) -> Optional[EvaluationResults]:
        return self.outcome.results if self.outcome else None

    def get_policy_score(self) -> float:
        return self.outcome.policy_score if self.outcome else 0.0
