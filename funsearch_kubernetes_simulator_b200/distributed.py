"""Multi-GPU plumbing: the episode batch shards across ranks, the only exchange is one all-gather of
per-episode fitness.  The reference's counterpart is the ProcessPoolExecutor that
maps policies over worker processes (funsearch/funsearch_integration.py:535-540) and collects
(index, code, score) tuples with as_completed.

Backend-agnostic torch.distributed code: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import Callable, List, Sequence

import numpy as np


def episode_shard(n_items: int, rank: int, world: int, costs: Sequence[float] = None) -> List[int]:
    """Indices of the items rank `rank` evaluates.  With `costs` (e.g. last generation's event counts) items are
    dealt longest-first to the currently lightest rank (LPT), which balances the ranks' kernel times;
    without, round-robin.  Deterministic and identical on every rank."""
    if costs is None:
        return list(range(rank, n_items, world))
    order = sorted(range(n_items), key=lambda i: (-float(costs[i]), i))
    load = [0.0] * world
    mine: List[int] = []
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        load[r] += float(costs[i])
        if r == rank:
            mine.append(i)
    return sorted(mine)


def gather_scores(local_scores: Sequence[float], local_indices: Sequence[int], n_items: int, device=None) -> List[float]:
    """All ranks contribute the scores of their shard; every rank gets the full list in item order.
    One all-gather of fixed-size (padded) fp64 blocks, as the kernel's results are fp64."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        out = [0.0] * n_items
        for i, s in zip(local_indices, local_scores):
            out[i] = float(s)
        return out
    cap = (n_items + world - 1) // world + 1
    dev = device if device is not None else "cpu"
    block = torch.full((2 * cap,), -1.0, dtype=torch.float64, device=dev)
    k = len(local_indices)
    if k > cap:
        raise ValueError("shard larger than the gather block")
    block[:k] = torch.as_tensor(np.asarray(local_indices, dtype=np.float64), device=dev)
    block[cap:cap + k] = torch.as_tensor(np.asarray(local_scores, dtype=np.float64), device=dev)
    gathered = torch.empty(world * 2 * cap, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(gathered, block)
    g = gathered.cpu().numpy().reshape(world, 2, cap)
    out = [0.0] * n_items
    seen = 0
    for r in range(world):
        for idx, sc in zip(g[r, 0], g[r, 1]):
            if idx >= 0:
                out[int(idx)] = float(sc)
                seen += 1
    if seen != n_items:
        raise RuntimeError(f"gathered {seen} scores for {n_items} items")
    return out


def evaluate_population_sharded(codes: Sequence[str], evaluate_local: Callable[[List[str]], List[float]],
                                costs: Sequence[float] = None, device=None) -> List[float]:
    """Fitness of a whole population on all ranks: shard, evaluate the local shard with `evaluate_local`
    (BatchEvaluator.evaluate_population on a GPU rank), all-gather."""
    import torch.distributed as dist
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    mine = episode_shard(len(codes), rank, world, costs)
    local = evaluate_local([codes[i] for i in mine]) if mine else []
    return gather_scores(local, mine, len(codes), device)
