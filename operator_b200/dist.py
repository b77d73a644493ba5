"""One-process-per-GPU plumbing (torch.distributed) for the benchmark and its CPU tests.

The path shards by pod with no data-path collective; the only
collectives are the barrier and the max/sum reductions of the measurement itself.
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def env_world() -> Tuple[int, int, int]:
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def init(backend: str) -> Tuple[int, int, int]:
    rank, local, world = env_world()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, local, world


def barrier():
    if dist.is_initialized():
        dist.barrier()


def _reduce(value: float, op, device: str) -> float:
    if not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=op)
    return float(t.item())


def max_over_ranks(value: float, device: str = "cpu") -> float:
    return _reduce(value, dist.ReduceOp.MAX, device)


def sum_over_ranks(value: float, device: str = "cpu") -> float:
    return _reduce(value, dist.ReduceOp.SUM, device)


def shard_pods(sizes: Sequence[int], world: int) -> List[List[int]]:
    """Longest-processing-time greedy partition of pods by bytes: the rule the C runtime applies
    across the devices of one context (pm_runtime.cu, stage()), restated for ranks."""
    order = sorted(range(len(sizes)), key=lambda i: (-sizes[i], i))
    load = [0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += sizes[i] + 64
    for lst in out:
        lst.sort()
    return out
