"""Limb/batch partition of an [L, B, N] workload across the GPUs of one box.

Every limb-polynomial is independent (twiddles depend only on (N, q_limb)), so the compute needs
no cross-GPU traffic.  Limbs are split first (each rank then holds only its limbs' tables), the
batch only when L < world.  The one collective is the final all-gather of result limbs back to
every caller's device (NCCL over NVLink; gloo in the CPU tests).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import torch


@dataclass(frozen=True)
class Shard:
    rank: int
    world: int
    g_l: int          # ranks along the limb axis
    g_b: int          # ranks along the batch axis
    limbs: Tuple[int, int]    # [lo, hi)
    batch: Tuple[int, int]    # [lo, hi)

    @property
    def num_limbs(self) -> int:
        return self.limbs[1] - self.limbs[0]

    @property
    def num_batch(self) -> int:
        return self.batch[1] - self.batch[0]


def partition(L: int, B: int, world: int, rank: int) -> Shard:
    """g_l = largest divisor of `world` that is <= L and divides L evenly when possible."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    g_l = 1
    for d in range(1, world + 1):
        if world % d == 0 and d <= L and L % d == 0:
            g_l = d
    g_b = world // g_l
    if B % g_b:
        raise ValueError("cannot split L=%d, B=%d evenly over %d ranks (g_l=%d, g_b=%d)" % (L, B, world, g_l, g_b))
    i, j = divmod(rank, g_b)
    lpr, bpr = L // g_l, B // g_b
    return Shard(rank, world, g_l, g_b, (i * lpr, (i + 1) * lpr), (j * bpr, (j + 1) * bpr))


def take(x: torch.Tensor, sh: Shard) -> torch.Tensor:
    """The rank's block of a full [L, B, N] tensor (contiguous copy)."""
    return x[sh.limbs[0]:sh.limbs[1], sh.batch[0]:sh.batch[1]].contiguous()


def all_gather_limbs(local: torch.Tensor, sh: Shard, group=None) -> torch.Tensor:
    """All-gather the per-rank [L/g_l, B/g_b, N] blocks into the full [L, B, N] on every rank."""
    import torch.distributed as dist
    l, b, n = local.shape
    flat = torch.empty((sh.world * l, b, n), dtype=local.dtype, device=local.device)   # concatenation form
    # torch.uint64 is not accepted by every backend: gather the bit pattern as int64
    dist.all_gather_into_tensor(flat.view(torch.int64), local.contiguous().view(torch.int64), group=group)
    if sh.g_b == 1:
        return flat.reshape(sh.g_l * l, b, n)       # limb-outermost layout: no repacking needed
    return flat.reshape(sh.g_l, sh.g_b, l, b, n).permute(0, 2, 1, 3, 4).reshape(sh.g_l * l, sh.g_b * b, n)


def bind_to_gpu_numa(device_index: int):
    """Pin the calling process to the CPU cores NVML reports as local to `device_index`, so that the pinned host buffers
    it allocates next (first-touch) and the threads that drive the copies sit on the GPU's own NUMA node.  Matters for
    the host-buffer entry points when several ranks share one box: without it the ranks' PCIe traffic crosses the
    socket interconnect.  Returns the previous affinity set (pass it to os.sched_setaffinity to undo), or None."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(int(device_index))
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {i for i in range(ncpu) if (int(words[i // 64]) >> (i % 64)) & 1}
        prev = os.sched_getaffinity(0)
        cpus &= prev
        if cpus and cpus != prev:
            os.sched_setaffinity(0, cpus)
            return prev
    except Exception:
        return None
    return None
