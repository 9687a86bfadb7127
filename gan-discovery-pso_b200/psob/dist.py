"""Multi-GPU plumbing: one process per GPU, particles sharded contiguously by rank, one
all-gather of (4 + D)-word candidate records per iteration (no other data-path collective).

record = [key_lo, key_hi, value_bits, 0, pos[0..D)] as 32-bit words, where
key = order_key(pbest_val) << 32 | global_particle_index; the minimum key over ranks is the
reference's in-order scan winner (lowest global index on ties, src/pso/util_pso.py:139-143).
"""
from __future__ import annotations

import os
from typing import Tuple

import numpy as np
import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block of particles owned by `rank`: (first global index, count)."""
    base, rem = divmod(total, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def order_key_np(x) -> np.ndarray:
    b = np.asarray(x, dtype=np.float32).view(np.uint32)
    b = np.where(b == np.uint32(0x80000000), np.uint32(0), b)          # -0.0 ties with +0.0 (csrc/psob_common.cuh order_key)
    return np.where(b >> 31 != 0, ~b, b | np.uint32(0x80000000)).astype(np.uint32)


def make_record(pbest_val: np.ndarray, pbest_pos: np.ndarray, first_index: int) -> np.ndarray:
    """Host mirror of psob_pbest_minloc + psob_gbest_candidate for one shard (protocol tests)."""
    keys = (order_key_np(pbest_val).astype(np.uint64) << np.uint64(32)) | (np.arange(len(pbest_val), dtype=np.uint64) + np.uint64(first_index))
    j = int(np.argmin(keys))
    rec = np.zeros(4 + pbest_pos.shape[1], dtype=np.float32)
    w = rec.view(np.uint32)
    w[0] = np.uint32(int(keys[j]) & 0xFFFFFFFF)
    w[1] = np.uint32(int(keys[j]) >> 32)
    rec[2] = pbest_val[j]
    rec[4:] = pbest_pos[j]
    return rec


def select_record(recs: np.ndarray) -> int:
    """Index of the winning record (minimum 64-bit key) — host mirror of psob_gbest_apply's choice."""
    w = np.ascontiguousarray(recs, dtype=np.float32).view(np.uint32)
    keys = (w[:, 1].astype(np.uint64) << np.uint64(32)) | w[:, 0].astype(np.uint64)
    return int(np.argmin(keys))


def init_from_env(backend: str = "nccl"):
    """torchrun-style initialisation; returns (rank, world, local_rank)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device(f"cuda:{local}"))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world, local
