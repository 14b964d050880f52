#!/usr/bin/env python
"""torchrun probe of the peer-store gather at config 4's shard: kernels alone, push alone, kernels + push, and the push
of chunk k under chunk k+1's kernels, for the CTA count in GFHE_PUSH_CTAS.  usage: torchrun ... tools/push_probe.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import polynomial as P  # noqa: E402
from gpu_fhe_b200 import _capi, shard  # noqa: E402

N, L, B = 65536, 32, 16
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lib = _capi.lib
sh = shard.partition(L, B, world, rank)
nl, nb = sh.num_limbs, sh.num_batch
moduli = P.ntt_primes(60, L)[sh.limbs[0]:sh.limbs[1]]
plan = P.make_plan(N, moduli, device=local)
gen = torch.Generator(device=dev); gen.manual_seed(rank)
nset = 4
da = [torch.randint(0, 1 << 59, (nl, nb, N), dtype=torch.int64, device=dev, generator=gen) for _ in range(nset)]
db = [torch.randint(0, 1 << 59, (nl, nb, N), dtype=torch.int64, device=dev, generator=gen) for _ in range(nset)]
wsb = lib.gfhe_poly_mul_workspace_bytes(plan.handle, nl, nb)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
pg = shard.PeerGather(L, B, N, dev)
mine = pg.local_block(sh)
main = torch.cuda.current_stream(dev)
st = ctypes.c_void_p(main.cuda_stream)
side = torch.cuda.Stream(dev)
limb_bytes = nb * N * 8


def kernels(i, k0=0, cl=None):
    cl = cl or nl
    o = k0 * limb_bytes
    rc = lib.gfhe_poly_mul(plan.handle, k0, cl, nb, da[i % nset].data_ptr() + o, db[i % nset].data_ptr() + o, mine.data_ptr() + o, ws.data_ptr(),
                           lib.gfhe_poly_mul_workspace_bytes(plan.handle, cl, nb), st)
    assert rc == 0, rc


def timed(f, steps=20):
    for i in range(3):
        f(i)
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    for i in range(steps):
        f(i)
    e1.record(main)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def both(i):
    kernels(i); pg.push(mine, sh); pg.finish()


def overlap(cl):
    def f(i):
        for k0 in range(0, nl, cl):
            kernels(i, k0, cl)
            ev = torch.cuda.Event(); ev.record(main); side.wait_event(ev)
            with torch.cuda.stream(side):
                for k in range(k0, k0 + cl):
                    pg.push_limb(mine, k, sh)
        ev = torch.cuda.Event(); ev.record(side); main.wait_event(ev)
        pg.finish()
    return f


res = {"kernels": timed(lambda i: kernels(i)), "push": timed(lambda i: (pg.push(mine, sh), pg.finish())), "kernels+push": timed(both)}
for cl in sorted({1, max(1, nl // 2)}):
    res["overlap cl=%d" % cl] = timed(overlap(cl))
if rank == 0:
    print("world %d  GFHE_PUSH_CTAS=%s  " % (world, os.environ.get("GFHE_PUSH_CTAS", "default")) + "  ".join("%s %.4f ms" % kv for kv in res.items()), flush=True)
pg.close()
dist.destroy_process_group()
