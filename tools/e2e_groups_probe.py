#!/usr/bin/env python
"""Host-buffer product (gfhe_poly_mul_host) of the config-4 shard of each of N ranks, for several pipeline group sizes
(GFHE_HOST_GROUP_MIB; below one limb's 8 MiB the library cuts the batch).  Wall time per call, max over ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 tools/e2e_groups_probe.py
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
import polynomial as P  # noqa: E402
from gpu_fhe_b200 import dispatch, shard  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
shard.bind_to_gpu_numa(local)
sh = shard.partition(bench.L_TOTAL, bench.B_TOTAL, world, rank)
all_moduli = P.ntt_primes(bench.BITS, bench.L_TOTAL)
moduli = all_moduli[sh.limbs[0]:sh.limbs[1]]
a_h, b_h = bench.shard_inputs(all_moduli, sh)
pa = torch.from_numpy(a_h.view(np.int64)).pin_memory().numpy().view(np.uint64)
pb = torch.from_numpy(b_h.view(np.int64)).pin_memory().numpy().view(np.uint64)
pc = torch.empty(a_h.shape, dtype=torch.int64).pin_memory().numpy().view(np.uint64)


def maxr(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


ref = None
for mib in [int(x) for x in (sys.argv[1:] or ["32", "8", "4", "2", "1", "32"])]:
    os.environ["GFHE_HOST_GROUP_MIB"] = str(mib)
    for _ in range(2):
        dispatch.poly_mul_host(pa, pb, pc, moduli, device=local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(12):
        dispatch.poly_mul_host(pa, pb, pc, moduli, device=local)
    own = (time.perf_counter() - t0) / 12
    s = maxr(own)
    hs = dispatch.host_stats(moduli, bench.N, device=local)
    if ref is None:
        ref = pc.copy()
    same = maxr(0.0 if np.array_equal(pc, ref) else 1.0) == 0.0
    slow = maxr(hs["h2d_ms"])
    if rank == 0:
        print("ranks=%d group=%2d MiB groups=%2d  %.3f ms per call (max over ranks; rank 0 own %.3f)  H2D busy max %.2f ms  same result: %s" % (
            world, mib, hs["groups"], s * 1e3, own * 1e3, slow, same), flush=True)
if world > 1:
    dist.destroy_process_group()
