import os, sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import polynomial as P
from gpu_fhe_b200 import dispatch
N, L, B = 65536, 32, 16
moduli = P.ntt_primes(60, L)
rng = np.random.default_rng(1)
a = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])
pa = torch.from_numpy(a.view(np.int64)).pin_memory(); pb = torch.from_numpy(a[::-1].copy().view(np.int64)).pin_memory(); pc = torch.empty_like(pa).pin_memory()
an, bn, cn = (t.numpy().view(np.uint64) for t in (pa, pb, pc))
for _ in range(3): dispatch.poly_mul_host(an, bn, cn, moduli)
t0 = time.perf_counter()
for _ in range(20): dispatch.poly_mul_host(an, bn, cn, moduli)
dt = (time.perf_counter() - t0) / 20
print("group MiB", os.environ.get("GFHE_HOST_GROUP_MIB", "32"), "ms/step %.3f" % (dt * 1e3), "H2D GB/s %.1f" % (2 * pa.numel() * 8 / dt / 1e9))
# raw PCIe rates for reference
d = torch.empty_like(pa, device='cuda')
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10): d.copy_(pa, non_blocking=True)
torch.cuda.synchronize(); print("raw H2D GB/s %.1f" % (10 * pa.numel() * 8 / (time.perf_counter() - t0) / 1e9))
