"""Which block of GPUs on this box has the fastest path to host memory?  Run as a SUBPROCESS (it creates a context on
every GPU, which the caller's process should not hold):

    python gpu-fhe_b200/_pcie_probe.py <gpus-per-block>

On the 8-GPU boxes of this pool the GPUs do not reach host memory equally (profiles/r02_run23_pcie_blocks.txt, H2D + D2H
at once, GB/s for the whole block): four at once, GPUs 0-3 move 103 and GPUs 4-7 198; in pairs {0,1} 105, {2,3} 103,
{4,5} 160, {6,7} 183; any single GPU 95-98.  A 4-rank job fed from host memory is therefore twice as fast on GPUs 4-7
(bench.py e2e 7.13 -> 4.21 ms).
The probe times every aligned block of `gpus-per-block` devices with all of its GPUs copying both ways at once and prints
one JSON line: {"devices": [...], "rates": {"0": GB/s, "4": GB/s}}.  Block 0 is kept unless another block is at least
10 % faster.  Only the CUDA runtime is used (ctypes); no torch import, no kernels.
"""
import ctypes
import json
import sys
import time

MIB = 1 << 20
BYTES = 64 * MIB
REPS = 6


def _cudart():
    for name in ("libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so", "libcudart.so"):
        try:
            return ctypes.CDLL(name)
        except OSError:
            continue
    import torch  # noqa: F401  (its wheel carries the runtime)
    return ctypes.CDLL("libcudart.so.12")


def _ck(rc, what):
    if rc:
        raise RuntimeError("%s: cuda error %d" % (what, rc))


class _Gpu:
    def __init__(self, rt, dev):
        self.rt, self.dev = rt, dev
        _ck(rt.cudaSetDevice(dev), "cudaSetDevice")
        self.h_in, self.h_out, self.d_in, self.d_out = (ctypes.c_void_p() for _ in range(4))
        _ck(rt.cudaHostAlloc(ctypes.byref(self.h_in), ctypes.c_size_t(BYTES), ctypes.c_uint(0)), "cudaHostAlloc")
        _ck(rt.cudaHostAlloc(ctypes.byref(self.h_out), ctypes.c_size_t(BYTES), ctypes.c_uint(0)), "cudaHostAlloc")
        _ck(rt.cudaMalloc(ctypes.byref(self.d_in), ctypes.c_size_t(BYTES)), "cudaMalloc")
        _ck(rt.cudaMalloc(ctypes.byref(self.d_out), ctypes.c_size_t(BYTES)), "cudaMalloc")
        ctypes.memset(self.h_in, 1, BYTES)
        ctypes.memset(self.h_out, 0, BYTES)
        self.s_in, self.s_out = ctypes.c_void_p(), ctypes.c_void_p()
        _ck(rt.cudaStreamCreate(ctypes.byref(self.s_in)), "cudaStreamCreate")
        _ck(rt.cudaStreamCreate(ctypes.byref(self.s_out)), "cudaStreamCreate")

    def issue(self):
        rt = self.rt
        _ck(rt.cudaSetDevice(self.dev), "cudaSetDevice")
        _ck(rt.cudaMemcpyAsync(self.d_in, self.h_in, ctypes.c_size_t(BYTES), 1, self.s_in), "H2D")
        _ck(rt.cudaMemcpyAsync(self.h_out, self.d_out, ctypes.c_size_t(BYTES), 2, self.s_out), "D2H")

    def wait(self):
        _ck(self.rt.cudaSetDevice(self.dev), "cudaSetDevice")
        _ck(self.rt.cudaDeviceSynchronize(), "cudaDeviceSynchronize")


def block_rate(gpus):
    """GB/s moved (both directions, all GPUs of the block at once)"""
    for g in gpus:          # warm-up
        g.issue()
    for g in gpus:
        g.wait()
    t0 = time.perf_counter()
    for _ in range(REPS):
        for g in gpus:
            g.issue()
    for g in gpus:
        g.wait()
    dt = time.perf_counter() - t0
    return 2.0 * BYTES * REPS * len(gpus) / dt / 1e9


def choose(per_block):
    rt = _cudart()
    n = ctypes.c_int(0)
    _ck(rt.cudaGetDeviceCount(ctypes.byref(n)), "cudaGetDeviceCount")
    n = n.value
    identity = list(range(per_block))
    if per_block < 1 or n < 2 * per_block:
        return {"devices": identity, "rates": {}}
    gpus = [_Gpu(rt, d) for d in range(n - n % per_block)]
    rates = {}
    for first in range(0, len(gpus), per_block):
        rates[str(first)] = round(block_rate(gpus[first:first + per_block]), 1)
    best = max(rates, key=lambda k: rates[k])
    first = int(best) if rates[best] >= 1.10 * rates["0"] else 0
    return {"devices": list(range(first, first + per_block)), "rates": rates}


def all_at_once():
    """--all: every GPU of the box copying both ways at once (the ceiling of a job that uses all of them)"""
    rt = _cudart()
    n = ctypes.c_int(0)
    _ck(rt.cudaGetDeviceCount(ctypes.byref(n)), "cudaGetDeviceCount")
    gpus = [_Gpu(rt, d) for d in range(n.value)]
    return {"devices": list(range(n.value)), "rates": {"all": round(block_rate(gpus), 1)}}


if __name__ == "__main__":
    print(json.dumps(all_at_once() if "--all" in sys.argv else choose(int(sys.argv[1]))), flush=True)
