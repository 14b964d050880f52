#!/usr/bin/env python
"""What bounds the host-buffer entry point on one GPU?  Config-4 operands in (a) torch-pinned memory, (b) pinned
write-combined memory for the two inputs (cudaHostAllocWriteCombined: no snooping on the DMA reads), with limb groups of
16 / 32 / 64 MiB, next to the raw copy rates of the same buffers.   usage: python tools/e2e_probe.py"""
import ctypes
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import polynomial as P  # noqa: E402
from gpu_fhe_b200 import dispatch  # noqa: E402

N, L, B = 65536, 32, 16
moduli = P.ntt_primes(60, L)
rng = np.random.default_rng(1)
a = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])
b = a[::-1].copy()
nbytes = a.nbytes
torch.cuda.init()
rt = ctypes.CDLL("libcudart.so.12")


def host_alloc(flags):
    p = ctypes.c_void_p()
    rc = rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(nbytes), ctypes.c_uint(flags))
    assert rc == 0, rc
    buf = (ctypes.c_uint64 * (nbytes // 8)).from_address(p.value)
    return np.frombuffer(buf, dtype=np.uint64).reshape(a.shape), p


def leg(tag, xa, xb, xc, group):
    os.environ["GFHE_HOST_GROUP_MIB"] = str(group)
    for _ in range(3):
        dispatch.poly_mul_host(xa, xb, xc, moduli)
    t0 = time.perf_counter()
    for _ in range(15):
        dispatch.poly_mul_host(xa, xb, xc, moduli)
    dt = (time.perf_counter() - t0) / 15
    hs = dispatch.host_stats(moduli, N)
    print("%-34s group %3d MiB  %.3f ms/step  (H2D busy %.2f ms = %.1f GB/s, D2H %.1f GB/s, kernels %.2f ms)" % (
        tag, group, dt * 1e3, hs["h2d_ms"], hs["h2d_bytes"] / hs["h2d_ms"] / 1e6, hs["d2h_bytes"] / hs["d2h_ms"] / 1e6, hs["kernel_ms"]), flush=True)


pa = torch.from_numpy(a.view(np.int64)).pin_memory().numpy().view(np.uint64)
pb = torch.from_numpy(b.view(np.int64)).pin_memory().numpy().view(np.uint64)
pc = torch.empty(a.shape, dtype=torch.int64).pin_memory().numpy().view(np.uint64)
wa, _ha = host_alloc(4)      # cudaHostAllocWriteCombined
wb, _hb = host_alloc(4)
t0 = time.perf_counter(); wa[...] = a; wb[...] = b; t_fill = time.perf_counter() - t0
print("filling the two write-combined inputs from pageable arrays: %.1f ms (%.1f GB/s)" % (t_fill * 1e3, 2 * nbytes / t_fill / 1e9))
for g in (16, 32, 64):
    leg("pinned (torch pin_memory)", pa, pb, pc, g)
want = pc.copy()
for g in (16, 32, 64):
    leg("inputs pinned write-combined", wa, wb, pc, g)
assert np.array_equal(pc, want)
# raw copy rates
d = torch.empty(a.shape, dtype=torch.int64, device="cuda")
for tag, src in (("pinned", pa), ("write-combined", wa)):
    t = torch.from_numpy(src.view(np.int64)) if tag == "pinned" else None
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10):
        rc = rt.cudaMemcpyAsync(ctypes.c_void_p(d.data_ptr()), ctypes.c_void_p(src.ctypes.data), ctypes.c_size_t(nbytes), 1, None)
        assert rc == 0
    torch.cuda.synchronize()
    print("raw H2D from %-15s %.1f GB/s" % (tag, 10 * nbytes / (time.perf_counter() - t0) / 1e9))
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10):
    rt.cudaMemcpyAsync(ctypes.c_void_p(pc.ctypes.data), ctypes.c_void_p(d.data_ptr()), ctypes.c_size_t(nbytes), 2, None)
torch.cuda.synchronize()
print("raw D2H to pinned             %.1f GB/s" % (10 * nbytes / (time.perf_counter() - t0) / 1e9))
