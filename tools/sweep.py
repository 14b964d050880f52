#!/usr/bin/env python
"""Throughput table over BASELINE.json's configs 2-5 on ONE GPU (device-resident operands, CUDA events on the launching
stream, inputs larger than L2 or rotated across enough buffers to exceed it).  Not the bench contract -- bench.py is;
this is the sweep DESIGN.md quotes.   usage: python tools/sweep.py [--quick] > gpurun_out/sweep.txt
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import polynomial as P  # noqa: E402
from gpu_fhe_b200 import _capi  # noqa: E402

L2_BYTES = 126e6


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6530.6


def run(op, N, L, B, bits, steps=30, warmup=5):
    dev = torch.device("cuda", 0)
    moduli = P.ntt_primes(bits, L)
    plan = P.make_plan(N, moduli, device=0)
    lib = _capi.lib
    packed = op.endswith("_p32")      # packed 32-bit layout (moduli below 2^30)
    op = op[:-4] if packed else op
    word = 4 if packed else 8
    bytes_per_buf = L * B * N * word
    # rotate over enough buffer sets that consecutive launches never find their inputs in L2
    nset = max(1, min(8, int(2 * L2_BYTES // bytes_per_buf) + 1))
    rng = np.random.default_rng(N + L + B)
    host = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli]).view(np.int64)
    if packed:
        host = host.astype(np.int32)
    a = [torch.from_numpy(host).to(dev) for _ in range(nset)]
    b = [torch.from_numpy(host[::-1].copy()).to(dev) for _ in range(nset)] if op == "polymul" else None
    c = [torch.empty_like(a[0]) for _ in range(nset)]
    ws_bytes = lib.gfhe_poly_mul_workspace_bytes(plan.handle, L, B) // (2 if packed else 1) if op == "polymul" else 0
    f_mul = lib.gfhe_poly_mul_u32 if packed else lib.gfhe_poly_mul
    f_fwd = lib.gfhe_ntt_forward_u32 if packed else lib.gfhe_ntt_forward
    f_inv = lib.gfhe_ntt_inverse_u32 if packed else lib.gfhe_ntt_inverse
    ws = torch.empty(max(ws_bytes, 8), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)
    st = ctypes.c_void_p(stream.cuda_stream)

    def step(i):
        k = i % nset
        if op == "polymul":
            rc = f_mul(plan.handle, 0, L, B, a[k].data_ptr(), b[k].data_ptr(), c[k].data_ptr(), ws.data_ptr(), ws_bytes, st)
        elif op == "ntt":
            rc = f_fwd(plan.handle, 0, L, B, a[k].data_ptr(), c[k].data_ptr(), st)
        else:
            rc = f_inv(plan.handle, 0, L, B, a[k].data_ptr(), c[k].data_ptr(), st)
        if rc:
            raise RuntimeError(_capi.error_string(rc))

    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(steps):
        step(i)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    units = L * B
    alg = (3 if op == "polymul" else 2) * word * N * units
    gbs = alg / (ms * 1e-3) / 1e9
    del a, b, c, ws
    P.clear_plans()
    torch.cuda.empty_cache()
    return {"op": op + ("_p32" if packed else ""), "N": N, "L": L, "B": B, "bits": bits, "ms": ms, "units_per_s": units / (ms * 1e-3), "alg_GBs": gbs,
            "frac_hbm": gbs / peak_gbs(), "buffer_sets": nset}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    cases = [
        ("polymul", 1024, 1, 1, 30), ("ntt", 1024, 1, 1, 30),                    # config 1 (launch-latency bound)
        ("polymul", 4096, 4, 256, 50),                                          # config 2
        ("ntt", 16384, 8, 128, 60), ("intt", 16384, 8, 128, 60),                # config 3
        ("polymul", 65536, 32, 16, 60), ("ntt", 65536, 32, 16, 60), ("intt", 65536, 32, 16, 60),   # config 4
    ]
    cases += [("polymul", 65536, 32, 16, 30), ("ntt", 65536, 32, 16, 30), ("intt", 65536, 32, 16, 30),   # 32-bit policy
              ("polymul", 4096, 4, 256, 30), ("ntt", 16384, 8, 128, 30),
              ("polymul_p32", 65536, 32, 16, 30), ("ntt_p32", 65536, 32, 16, 30), ("intt_p32", 65536, 32, 16, 30),   # packed layout
              ("polymul_p32", 4096, 4, 256, 30), ("ntt_p32", 16384, 8, 128, 30), ("ntt_p32", 131072, 16, 16, 30),
              ("polymul", 65536, 32, 16, 50), ("ntt", 65536, 32, 16, 50), ("intt", 65536, 32, 16, 50), ("ntt", 16384, 8, 128, 50)]   # FP64 policy
    grid = [1, 4, 16, 64] if not args.quick else [1, 16]
    for L in grid:                                                              # config 5: N = 2^17 sweep
        for B in grid:
            cases.append(("ntt", 131072, L, B, 60))
    print("%-12s %7s %3s %4s %4s %10s %14s %10s %8s" % ("op", "N", "L", "B", "bits", "ms/launch", "limb-ops/s", "alg GB/s", "% HBM"))
    out = []
    for op, N, L, B, bits in cases:
        r = run(op, N, L, B, bits)
        out.append(r)
        print("%-12s %7d %3d %4d %4d %10.4f %14.0f %10.1f %7.1f%%" % (op, N, L, B, bits, r["ms"], r["units_per_s"], r["alg_GBs"],
                                                                 100 * r["frac_hbm"]), flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
