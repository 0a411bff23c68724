"""Multi-GPU plumbing: the episode batch shards across ranks, the only exchange is one all-gather of
per-episode fitness.  The reference's counterpart is the ProcessPoolExecutor that
maps policies over worker processes (funsearch/funsearch_integration.py:535-540) and collects
(index, code, score) tuples with as_completed.

Backend-agnostic torch.distributed code: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import Callable, List, Sequence

import numpy as np


def shard_plan(n_items: int, world: int, costs: Sequence[float] = None) -> List[List[int]]:
    """The items of every rank.  With `costs` (e.g. last generation's event counts) items are dealt longest-first to
    the currently lightest rank (LPT), which balances the ranks' kernel times; zero-cost items (last generation's
    failed policies) go to the rank holding the fewest items instead of piling onto the lightest one.
    Without costs: round-robin.  Deterministic, and every rank computes the same plan."""
    if costs is None:
        return [list(range(r, n_items, world)) for r in range(world)]
    if len(costs) != n_items:
        raise ValueError("one cost per item")
    order = sorted(range(n_items), key=lambda i: (-float(costs[i]), i))
    load = [0.0] * world
    shards: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        c = max(0.0, float(costs[i]))
        if c > 0.0:
            r = min(range(world), key=lambda k: (load[k], len(shards[k]), k))
        else:   # free items change no load: even out the item counts with them
            r = min(range(world), key=lambda k: (len(shards[k]), load[k], k))
        load[r] += c
        shards[r].append(i)
    return [sorted(sh) for sh in shards]


def episode_shard(n_items: int, rank: int, world: int, costs: Sequence[float] = None) -> List[int]:
    """Indices of the items rank `rank` evaluates (see shard_plan)."""
    return shard_plan(n_items, world, costs)[rank]


def gather_scores(local_scores: Sequence[float], local_indices: Sequence[int], n_items: int, device=None,
                  plan: List[List[int]] = None) -> List[float]:
    """All ranks contribute the scores of their shard; every rank gets the full list in item order.
    One all-gather of fixed-size fp64 blocks, as the kernel's results are fp64.  The block holds (index, score)
    pairs and is as long as the LARGEST shard: with `plan` (shard_plan's result) that is known exactly on every
    rank; without it the sizes are agreed on by a first, tiny all-gather.  No rank raises before a collective the
    others have already entered."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size() if dist.is_initialized() else 1
    if len(local_scores) != len(local_indices):
        raise ValueError("one score per local index")   # a local programming error: every rank sees it or none
    if world == 1:
        out = [0.0] * n_items
        for i, s in zip(local_indices, local_scores):
            out[i] = float(s)
        return out
    dev = device if device is not None else "cpu"
    k = len(local_indices)
    if plan is not None:
        cap = max(len(sh) for sh in plan)
    else:
        sizes = torch.zeros(world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(sizes, torch.tensor([k], dtype=torch.int64, device=dev))
        cap = int(sizes.max().item())
    cap = max(cap, 1)
    block = torch.full((2 * cap,), -1.0, dtype=torch.float64, device=dev)
    fits = k <= cap   # always true when every rank used the same plan; checked after the collective
    if fits and k:
        block[:k] = torch.as_tensor(np.asarray(local_indices, dtype=np.float64), device=dev)
        block[cap:cap + k] = torch.as_tensor(np.asarray(local_scores, dtype=np.float64), device=dev)
    gathered = torch.empty(world * 2 * cap, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(gathered, block)
    if not fits:
        raise ValueError(f"local shard ({k} items) does not match the sharding plan (largest shard {cap})")
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
    plan = shard_plan(len(codes), world, costs)
    mine = plan[rank]
    local = evaluate_local([codes[i] for i in mine]) if mine else []
    return gather_scores(local, mine, len(codes), device, plan)
