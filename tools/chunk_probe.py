#!/usr/bin/env python
"""Does an L2-sized chunk schedule beat one launch per pass?  Config-4 shape [32,16,65536]; the limb range is split into
chunks of `cl` limbs, chunk k runs (column pass, row pass[, inverse column pass]) on stream k % ns, so a chunk's
intermediate is re-read while it is still in the 126 MB L2 and the other streams' kernels fill the launch tails.
Forked from / joined to the timing stream with events.  usage: python tools/chunk_probe.py [bits ...]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import polynomial as P  # noqa: E402
from gpu_fhe_b200 import _capi  # noqa: E402

N, L, B = 65536, int(os.environ.get("PROBE_L", "32")), 16
lib = _capi.lib
dev = torch.device("cuda", 0)


def run(bits, packed):
    moduli = P.ntt_primes(bits, L)
    plan = P.make_plan(N, moduli, device=0)
    rng = np.random.default_rng(0)
    h = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli]).view(np.int64)
    if packed:
        h = h.astype(np.int32)
    a = torch.from_numpy(h).to(dev)
    b = torch.from_numpy(h[::-1].copy()).to(dev)
    c = torch.empty_like(a)
    word = 4 if packed else 8
    ws = torch.empty(L * B * N * word, dtype=torch.uint8, device=dev)
    f_mul, f_fwd = (lib.gfhe_poly_mul_u32, lib.gfhe_ntt_forward_u32) if packed else (lib.gfhe_poly_mul, lib.gfhe_ntt_forward)
    main = torch.cuda.current_stream(dev)
    streams = [torch.cuda.Stream(dev) for _ in range(3)]
    limb_words = B * N

    def step(op, cl, ns):
        if ns == 0:          # everything on the timing stream, one call
            st = ctypes.c_void_p(main.cuda_stream)
            if op == "ntt":
                rc = f_fwd(plan.handle, 0, L, B, a.data_ptr(), c.data_ptr(), st)
            else:
                rc = f_mul(plan.handle, 0, L, B, a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), ws.numel(), st)
            assert rc == 0
            return
        ev = torch.cuda.Event()
        ev.record(main)
        used = streams[:ns]
        for s in used:
            s.wait_event(ev)
        for k, l0 in enumerate(range(0, L, cl)):
            s = used[k % ns]
            st = ctypes.c_void_p(s.cuda_stream)
            off = l0 * limb_words * word
            if op == "ntt":
                rc = f_fwd(plan.handle, l0, cl, B, a.data_ptr() + off, c.data_ptr() + off, st)
            else:
                rc = f_mul(plan.handle, l0, cl, B, a.data_ptr() + off, b.data_ptr() + off, c.data_ptr() + off, ws.data_ptr() + off,
                           cl * limb_words * word, st)
            assert rc == 0
        for s in used:
            e = torch.cuda.Event()
            e.record(s)
            main.wait_event(e)

    ref = {}
    for op in ("ntt", "polymul"):
        step(op, L, 0)
        torch.cuda.synchronize()
        ref[op] = c.clone()
        for cl, ns in [(cl, ns) for cl, ns in [(L, 0), (8, 1), (4, 1), (2, 1), (8, 2), (4, 2), (2, 2), (1, 2), (4, 3), (2, 3), (1, 3)] if cl <= L and L % cl == 0 and (ns == 0 or cl < L)]:
            for _ in range(3):
                step(op, cl, ns)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main)
            for _ in range(20):
                step(op, cl, ns)
            e1.record(main)
            torch.cuda.synchronize()
            ok = torch.equal(c, ref[op])
            print("bits=%d%s %-7s chunk=%2d limbs (%5.1f MiB/operand) streams=%d  %.4f ms  %s" % (
                bits, " packed" if packed else "", op, cl, cl * limb_words * word / 2**20, ns, e0.elapsed_time(e1) / 20, "ok" if ok else "MISMATCH"), flush=True)
    P.clear_plans()


if __name__ == "__main__":
    which = [int(x) for x in sys.argv[1:]] or [60, 50, 30]
    for bits in which:
        run(bits, False)
    if 30 in which:
        run(30, True)
