"""Limb/batch partition of an [L, B, N] workload across the GPUs of one box.

Every limb-polynomial is independent (twiddles depend only on (N, q_limb)), so the compute needs
no cross-GPU traffic.  Limbs are split first (each rank then holds only its limbs' tables), the
batch only when L < world.  The one collective is the final all-gather of result limbs back to
every caller's device (NCCL over NVLink; gloo in the CPU tests).

Two limb-to-rank maps.  Block (default): rank i holds limbs [i*L/g, (i+1)*L/g) -- one all-gather of
the whole shard lands in natural limb order.  Cyclic: rank i holds limbs i, i+g, i+2g, ... -- then
the k-th local limbs of all ranks are the CONTIGUOUS slab [k*g, (k+1)*g) of the result, so the
gather can be issued limb by limb, each one overlapping the next limb's kernels
(all_gather_limb_async); with the block map a per-limb gather would land strided.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch


@dataclass(frozen=True)
class Shard:
    rank: int
    world: int
    g_l: int          # ranks along the limb axis
    g_b: int          # ranks along the batch axis
    limbs: Tuple[int, int]    # block map: [lo, hi);  cyclic map: (first limb, L) -- use limb_ids
    batch: Tuple[int, int]    # [lo, hi)
    cyclic: bool = False

    @property
    def limb_ids(self) -> Tuple[int, ...]:
        """Global indices of the limbs this rank holds, in local order."""
        if self.cyclic:
            return tuple(range(self.limbs[0], self.limbs[1], self.g_l))
        return tuple(range(self.limbs[0], self.limbs[1]))

    @property
    def num_limbs(self) -> int:
        return len(self.limb_ids)

    @property
    def num_batch(self) -> int:
        return self.batch[1] - self.batch[0]


def partition(L: int, B: int, world: int, rank: int, cyclic: bool = False) -> Shard:
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
    if cyclic:
        return Shard(rank, world, g_l, g_b, (i, L), (j * bpr, (j + 1) * bpr), True)
    return Shard(rank, world, g_l, g_b, (i * lpr, (i + 1) * lpr), (j * bpr, (j + 1) * bpr))


def take(x: torch.Tensor, sh: Shard) -> torch.Tensor:
    """The rank's block of a full [L, B, N] tensor (contiguous copy)."""
    if sh.cyclic:
        return x[sh.limbs[0]:sh.limbs[1]:sh.g_l, sh.batch[0]:sh.batch[1]].contiguous()
    return x[sh.limbs[0]:sh.limbs[1], sh.batch[0]:sh.batch[1]].contiguous()


def all_gather_limbs(local: torch.Tensor, sh: Shard, group=None) -> torch.Tensor:
    """All-gather the per-rank [L/g_l, B/g_b, N] blocks into the full [L, B, N] on every rank."""
    import torch.distributed as dist
    l, b, n = local.shape
    if sh.cyclic:                      # limb by limb, each into its contiguous slab of the result
        full = torch.empty((sh.g_l * l, sh.g_b * b, n), dtype=local.dtype, device=local.device)
        works = [all_gather_limb_async(local[k], full, k, sh, group) for k in range(l)]
        for w in works:
            w.wait()
        return full
    flat = torch.empty((sh.world * l, b, n), dtype=local.dtype, device=local.device)   # concatenation form
    # torch.uint64 is not accepted by every backend: gather the bit pattern as int64
    dist.all_gather_into_tensor(flat.view(torch.int64), local.contiguous().view(torch.int64), group=group)
    if sh.g_b == 1:
        return flat.reshape(sh.g_l * l, b, n)       # limb-outermost layout: no repacking needed
    return flat.reshape(sh.g_l, sh.g_b, l, b, n).permute(0, 2, 1, 3, 4).reshape(sh.g_l * l, sh.g_b * b, n)


def all_gather_limb_async(local_k: torch.Tensor, full: torch.Tensor, k: int, sh: Shard, group=None):
    """Cyclic map only: start the all-gather of every rank's k-th local limb ([B/g_b, N]) into full[k*g_l:(k+1)*g_l] and
    return the work handle.  The collective is ordered after what the current stream has enqueued so far (the kernels that
    produce local_k) and runs on the backend's own stream, so kernels launched next -- limb k+1 -- overlap it; call
    .wait() on the handle before the current stream reads `full`.  Rank (i, j) of the (g_l x g_b) grid contributes limb
    k*g_l + i, batch block j: chunk i*g_b + j = its rank, which is exactly where that block sits in the slab."""
    import torch.distributed as dist
    if not sh.cyclic:
        raise ValueError("per-limb gathers land contiguously only with the cyclic limb map (partition(..., cyclic=True))")
    slab = full[k * sh.g_l:(k + 1) * sh.g_l]
    return dist.all_gather_into_tensor(slab.view(torch.int64).view(-1), local_k.contiguous().view(torch.int64).view(-1), group=group, async_op=True)


class PeerGather:
    """The all-gather of result limbs as NVLink peer stores issued by the library (csrc/gather.cu, gfhe_push_gather) instead of
    an NCCL all-gather after the kernels.

    Every rank owns a full [L, B, N] result array in peer-mappable memory (`.full`, a torch view of it); the ranks exchange
    CUDA IPC handles once (all_gather_object) and map each other's arrays.  `push(local, sh)` copies this rank's block --
    or, limb by limb, `push_limb(local, k, sh)` -- into the same position of every rank's array, asynchronously on the
    current stream (enqueue it behind the kernels of limb k and limb k+1's kernels overlap it).  `finish()` is the one point
    of cross-rank synchronisation: a stream-ordered all-reduce of a flag, which every rank enters after its own pushes, so
    when it completes on a rank every peer's block has landed there.  No kernel waits for another GPU.
    """

    def __init__(self, L: int, B: int, N: int, device: torch.device, group=None, word_bytes: int = 8):
        import ctypes
        import torch.distributed as dist
        from . import _capi
        self._capi, self._ct, self._dist = _capi, ctypes, dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > 16:
            raise ValueError("PeerGather serves the GPUs of one box (at most 16)")
        self.dev = torch.device(device)
        self.shape, self.word = (int(L), int(B), int(N)), int(word_bytes)
        self.bytes = L * B * N * word_bytes
        own = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        _capi.check(_capi.lib.gfhe_peer_alloc(self.dev.index, self.bytes, ctypes.byref(own), handle), "gfhe_peer_alloc")
        self._own = own
        handles = [None] * self.world
        dist.all_gather_object(handles, handle.raw, group=group)
        self._peers = []
        bases = []
        for r in range(self.world):
            if r == self.rank:
                bases.append(own.value)
                continue
            p = ctypes.c_void_p()
            _capi.check(_capi.lib.gfhe_peer_open(self.dev.index, handles[r], ctypes.byref(p)), "gfhe_peer_open")
            self._peers.append(p)
            bases.append(p.value)
        self._bases = (ctypes.c_void_p * self.world)(*bases)
        others = [b for r, b in enumerate(bases) if r != self.rank]
        self._others = (ctypes.c_void_p * max(1, len(others)))(*others)
        self.full = self._view(own.value)
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.dev)

    def _view(self, ptr: int) -> torch.Tensor:
        dt = torch.int64 if self.word == 8 else torch.int32
        typestr = "<i8" if self.word == 8 else "<i4"

        class _Mem:      # CUDA array interface over the raw allocation
            __cuda_array_interface__ = {"shape": self.shape, "typestr": typestr, "data": (ptr, False), "version": 3}
        t = torch.as_tensor(_Mem(), device=self.dev)
        assert t.dtype == dt and t.data_ptr() == ptr
        return t

    def local_block(self, sh: "Shard") -> Optional[torch.Tensor]:
        """This rank's block as a view INTO its own full array, when the block is contiguous there (block limb map, no
        batch split): have the kernels write their result here and the push need not copy it to this rank itself."""
        if sh.cyclic or sh.g_b != 1:
            return None
        return self.full[sh.limbs[0]:sh.limbs[1]]

    def _in_place(self, src: torch.Tensor, offset_bytes: int) -> bool:
        return src.data_ptr() == self._own.value + offset_bytes

    def _push(self, src: torch.Tensor, offset_bytes: int, ctas: int = 0) -> None:
        st = self._ct.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)
        # a block that already sits in this rank's array (local_block) goes to the peers only
        bases, n = (self._others, self.world - 1) if self._in_place(src, offset_bytes) else (self._bases, self.world)
        if n == 0:
            return
        self._capi.check(self._capi.lib.gfhe_push_gather(self.dev.index, self._ct.c_void_p(src.data_ptr()), src.numel() * src.element_size(),
                                                         bases, n, offset_bytes, int(ctas), st), "gfhe_push_gather")

    OVERLAP_CTAS = 32      # copy CTAs of a push that runs under other kernels (measured: csrc/gather.cu)

    def push_limb(self, local: torch.Tensor, k: int, sh: "Shard", ctas: Optional[int] = None) -> None:
        """Local limb k ([B/g_b, N], contiguous) of this rank's result into its place in every rank's array.  Meant to be
        enqueued on a side stream behind limb k's kernels, so by default it runs on few CTAs."""
        L, B, N = self.shape
        limb = sh.limb_ids[k]
        self._push(local[k], ((limb * B + sh.batch[0]) * N) * self.word, self.OVERLAP_CTAS if ctas is None else ctas)

    def push(self, local: torch.Tensor, sh: "Shard") -> None:
        """The whole local block [L/g_l, B/g_b, N]: one copy when it is contiguous in the result, else limb by limb."""
        L, B, N = self.shape
        if not sh.cyclic and sh.g_b == 1:
            self._push(local, (sh.limbs[0] * B * N) * self.word)
        else:
            for k in range(local.shape[0]):
                self.push_limb(local, k, sh, ctas=0)

    def finish(self) -> None:
        """Stream-ordered: when this returns control to the stream, every rank's pushes issued before its own finish() have landed."""
        self._dist.all_reduce(self._flag, group=self.group)

    def close(self) -> None:
        """Collective: every rank of the group must call it (peers unmap each other's arrays before any of them is freed)."""
        if getattr(self, "_own", None) is None:
            return
        torch.cuda.synchronize(self.dev)
        self._dist.barrier(group=self.group)          # nobody unmaps or frees while a peer may still be pushing
        for p in self._peers:
            self._capi.lib.gfhe_peer_close(self.dev.index, p)
        self._peers = []
        self._dist.barrier(group=self.group)
        self.full = None
        self._capi.lib.gfhe_peer_free(self.dev.index, self._own)
        self._own = None


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


def choose_local_devices(world: int, local_rank: int, timeout_s: float = 120.0):
    """Which GPU should the rank with this `local_rank` use when the box shows more GPUs than the job has ranks?

    On this pool's 8-GPU boxes GPUs 0-3 together reach half the host-memory bandwidth of GPUs 4-7 (`_pcie_probe.py`), so a
    4-rank job that feeds its GPUs from host memory (`poly_mul_host`) runs faster on 4-7.  Local rank 0 times the aligned
    blocks of `world` GPUs in a subprocess (no CUDA context is created in THIS process) and leaves the answer in a file
    named after MASTER_PORT in the temporary directory; the other ranks read it.  Local rank 0 always writes an answer (the
    identity map when the probe fails or times out), so the ranks cannot disagree.

    Opt in with GFHE_DEVICE_MAP=probe; anything else (the default) keeps device = local_rank.  `world` is the number of
    ranks on THIS box.  Returns (device_index, probe_result or None).
    """
    import json
    import os
    import subprocess
    import sys
    import tempfile
    import time

    if os.environ.get("GFHE_DEVICE_MAP", "identity") != "probe" or world <= 1:
        return local_rank, None
    try:
        if torch.cuda.device_count() < 2 * world:
            return local_rank, None
    except Exception:
        return local_rank, None
    path = os.path.join(tempfile.gettempdir(), "gfhe_devmap_%s.json" % os.environ.get("MASTER_PORT", "0"))
    if local_rank == 0:
        try:
            os.remove(path)                       # a crashed earlier job's answer
        except OSError:
            pass
        res = {"devices": list(range(world)), "rates": {}, "error": None}
        try:
            probe = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_pcie_probe.py")
            out = subprocess.run([sys.executable, probe, str(world)], capture_output=True, text=True, timeout=timeout_s)
            if out.returncode != 0:
                raise RuntimeError((out.stderr.strip().splitlines() or ["exit %d" % out.returncode])[-1])
            got = json.loads(out.stdout.strip().splitlines()[-1])
            if sorted(got["devices"]) == got["devices"] and len(got["devices"]) == world:
                res.update(got)
        except Exception as e:   # probe failed: identity, and say why
            res["error"] = "%s: %s" % (type(e).__name__, e)
        tmp = path + ".%d.tmp" % os.getpid()
        with open(tmp, "w") as f:
            json.dump(res, f)
        os.replace(tmp, path)
        return int(res["devices"][0]), res
    # the ranks of one job start within a moment of each other and local rank 0 writes its answer after that; a file older
    # than this process (less a margin) is a crashed earlier job's on the same port
    deadline = time.time() + timeout_s + 60.0
    born = _process_start_time(os.getpid())
    while time.time() < deadline:
        try:
            if os.stat(path).st_mtime >= born - 2.0:
                with open(path) as f:
                    res = json.load(f)
                if len(res["devices"]) == world:
                    return int(res["devices"][local_rank]), res
        except (OSError, ValueError, KeyError):
            pass
        time.sleep(0.05)
    raise RuntimeError("choose_local_devices: local rank 0 left no device map in %s" % path)


def _process_start_time(pid: int) -> float:
    """Unix time at which process `pid` started (0.0 when /proc does not say)"""
    import os
    import time
    try:
        with open("/proc/%d/stat" % pid) as f:
            ticks = int(f.read().rsplit(")", 1)[1].split()[19])      # field 22, starttime, after the "(comm)" field
        with open("/proc/uptime") as f:
            up = float(f.read().split()[0])
        return time.time() - (up - ticks / os.sysconf("SC_CLK_TCK"))
    except (OSError, ValueError, IndexError):
        return 0.0
