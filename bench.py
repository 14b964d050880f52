#!/usr/bin/env python
"""Benchmark of the NTT / poly-mul hot path (BASELINE.json metric: NTT/s and poly-mul/s at N=2^16).

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload ("config 4"): negacyclic poly-mul at N=2^16 over 32 RNS limbs of 60-bit primes, batch 16
per GPU -- 512 limb-poly-muls per GPU per step, 256 MiB per operand (inputs + output + scratch
= 1 GiB per step, far larger than the 126 MB L2, so no L2 flush is needed between iterations).
At n GPUs the batch is 16*n and the limbs are sharded (32/n limbs x 16*n batch per rank): per-GPU
work is fixed ("weak" scaling), compute needs no collective; the final all-gather is timed apart.

One JSON line on stdout (rank 0).  value = limb-poly-muls/s with operands resident in HBM, timed
with CUDA events on the launching stream, max over ranks.  e2e = the same through the host-buffer
C-ABI call (pinned host arrays, H2D + kernels + D2H inside the timed region).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOG_N = 16
N = 1 << LOG_N
L_TOTAL = 32
B_PER_GPU = 16
BITS = 60
UNITS_PER_GPU = L_TOTAL * B_PER_GPU            # limb-poly-muls per GPU per step
ALG_BYTES_POLYMUL = 24 * N                     # per limb-poly-mul: read a, read b, write c
ALG_BYTES_NTT = 16 * N                         # per limb-NTT
FALLBACK_HBM_GBS = 6650.0                      # B200_PROFILING.md fallback


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """dram bytes per launch of `kernel` from the committed ncu summary, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f).get(kernel)
    except Exception:
        return None


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while `active` (NVML, ~every 5 ms)."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, torch_index):
        self.samples, self.reasons, self.active, self.stop = [], set(), False, False
        self.max_mhz = None
        self.ok = False
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(torch_index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(torch_index)
            self.nv = pynvml
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
            self.t = threading.Thread(target=self._run, daemon=True)
            self.t.start()
        except Exception as e:   # pragma: no cover
            log("clock sampler unavailable:", e)

    def _run(self):
        nv = self.nv
        while not self.stop:
            if self.active:
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    try:
                        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                    except Exception:
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, name in self.REASONS.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
            time.sleep(0.004)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


def make_inputs(moduli, B, seed):
    import numpy as np
    rng = np.random.default_rng(seed)
    a = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])
    b = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])
    return a, b


def cpu_baseline(moduli, threads, reps):
    """C oracle (a CPU port -- the reference has no NTT), fast Shoup/Harvey flavour, all given threads."""
    import numpy as np
    from oracle import cref
    from oracle import oracle as O
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    a, b = make_inputs(moduli, B_PER_GPU, 1)
    R.poly_mul(a[:, :1], b[:, :1], fast=True, nthreads=threads)     # warm-up (page in, spin up threads)
    t0 = time.perf_counter()
    for _ in range(reps):
        c = R.poly_mul(a, b, fast=True, nthreads=threads)
    dt = time.perf_counter() - t0
    assert c.shape == a.shape and np.all(c[0, 0] < moduli[0])
    units = len(moduli) * B_PER_GPU * reps
    return units / dt, dt / reps, units


def run_reference(args, rank, world):
    """Reference arm: the path's CPU implementation on the box's host cores.  The reference repo
    has no NTT/poly-mul (identity stubs), so this is the oracle port (kind = "port")."""
    if rank != 0:
        return
    from oracle import cref
    from oracle import oracle as O
    cref.build()
    threads = os.cpu_count() or 1
    moduli = O.ntt_primes(BITS, L_TOTAL)
    import numpy as np
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    a, b = make_inputs(moduli, B_PER_GPU, 1)
    for _ in range(max(args.warmup, 1)):
        R.poly_mul(a[:, :2], b[:, :2], fast=True, nthreads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        c = R.poly_mul(a, b, fast=True, nthreads=threads)
    dt = time.perf_counter() - t0
    assert np.all(c[0, 0] < moduli[0])
    value = UNITS_PER_GPU * args.steps / dt
    sample = "each step = one full config-4 batch [32,16,65536] (512 limb-poly-muls) on %d host threads" % threads
    line = {
        "impl": "reference", "metric": "limb_poly_mul_per_s", "value": value, "unit": "limb-poly-mul/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": value, "unit": "limb-poly-mul/s", "cores": threads, "kind": "port", "sample": sample,
                         "note": "reference repo has no NTT/poly-mul (arithmetic.py:15-19 are identity stubs); "
                                 "this is oracle/ntt_oracle.c, Shoup/Harvey flavour, pthreads over limb-polynomials"},
        "e2e": {"value": value, "unit": "limb-poly-mul/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus):
    return {"workload": "config 4: negacyclic poly-mul, N=65536, 32 limbs of 60-bit primes, batch 16 per GPU "
                        "(global [32, %d, 65536], limb-sharded %d limbs x %d batch per GPU)"
                        % (B_PER_GPU * n_gpus, L_TOTAL // n_gpus, B_PER_GPU * n_gpus),
            "N": N, "limbs": L_TOTAL, "batch_per_gpu": B_PER_GPU, "modulus_bits": BITS,
            "units_per_gpu_per_step": UNITS_PER_GPU, "parallelism": "limb-shard x%d" % n_gpus,
            "l2_policy": "inputs larger than L2 (1 GiB touched per step vs 126 MB L2); no flush"}


def run_b200(args, rank, world, local_rank):
    import ctypes

    import numpy as np
    import torch
    import torch.distributed as dist

    import polynomial as P
    from gpu_fhe_b200 import _capi, dispatch, shard

    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # several ranks on one box: keep each rank's pinned buffers and copy threads on its GPU's NUMA node (e2e leg)
    prev_affinity = shard.bind_to_gpu_numa(local_rank) if world > 1 and not os.environ.get("GFHE_BENCH_NO_NUMA") else None
    B = B_PER_GPU * world
    sh = shard.partition(L_TOTAL, B, world, rank)
    assert sh.g_b == 1 and sh.num_limbs * sh.num_batch == UNITS_PER_GPU
    all_moduli = P.ntt_primes(BITS, L_TOTAL)
    moduli = all_moduli[sh.limbs[0]:sh.limbs[1]]
    nl, nb = sh.num_limbs, sh.num_batch
    plan = P.make_plan(N, moduli, device=local_rank)
    lib = _capi.lib

    a_h, b_h = make_inputs(moduli, nb, 20261104 + 4 + 1000 * rank)
    pa = torch.from_numpy(a_h.view(np.int64)).pin_memory()
    pb = torch.from_numpy(b_h.view(np.int64)).pin_memory()
    pc = torch.empty_like(pa).pin_memory()
    da, db = pa.to(dev), pb.to(dev)
    dc = torch.empty_like(da)
    ws_bytes = lib.gfhe_poly_mul_workspace_bytes(plan.handle, nl, nb)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)
    st = ctypes.c_void_p(stream.cuda_stream)

    def step_polymul():
        rc = lib.gfhe_poly_mul(plan.handle, 0, nl, nb, da.data_ptr(), db.data_ptr(), dc.data_ptr(), ws.data_ptr(),
                               ws_bytes, st)
        if rc:
            raise RuntimeError(_capi.error_string(rc))

    def step_ntt():
        rc = lib.gfhe_ntt_forward(plan.handle, 0, nl, nb, da.data_ptr(), dc.data_ptr(), st)
        if rc:
            raise RuntimeError(_capi.error_string(rc))

    def timed(step, steps, warmup):
        for _ in range(warmup):
            step()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        barrier()
        return max_over_ranks(ms)

    clocks = ClockSampler(local_rank)
    K, W = args.steps, max(args.warmup, 3)

    # ---- correctness guard: one limb-polynomial of the result against the oracle (checker only)
    step_polymul()
    torch.cuda.synchronize()
    if rank == 0:
        from oracle import cref
        from oracle import oracle as O
        R1 = cref.Ring(N, moduli[:1], [O.find_psi(moduli[0], N)])
        want = R1.poly_mul(a_h[:1, :1], b_h[:1, :1], fast=True)
        got = dc[0, 0].cpu().numpy().view(np.uint64)
        assert np.array_equal(got, want[0, 0]), "bench result is not bit-exact"

    # ---- headline: device-resident poly-mul
    clocks.active = True
    ms_total = timed(step_polymul, K, W)
    # ---- per-kernel live durations (events around every launch, same loop)
    lib.gfhe_profile_enable(plan.handle, 1)
    timed(step_polymul, K, 1)
    names = (ctypes.c_char_p * 8)()
    tot = (ctypes.c_double * 8)()
    cnt = (ctypes.c_uint64 * 8)()
    nk = lib.gfhe_profile_read(plan.handle, 8, names, tot, cnt)
    lib.gfhe_profile_enable(plan.handle, 0)
    kernels = {names[i].decode(): {"ms_per_launch": tot[i] / cnt[i], "launches_per_step": cnt[i] / (K + 1)} for i in range(max(nk, 0))}
    # ---- forward NTT throughput (second half of the metric)
    ms_ntt = timed(step_ntt, K, W)

    def step_intt():
        rc = lib.gfhe_ntt_inverse(plan.handle, 0, nl, nb, da.data_ptr(), dc.data_ptr(), st)
        if rc:
            raise RuntimeError(_capi.error_string(rc))
    ms_intt = timed(step_intt, K, W)
    # ---- NTT-domain-resident kernels: pure HBM streaming, timed the same way
    inner = nb * N
    da4 = da.unsqueeze(0).expand(4, nl, nb, N).contiguous()
    db4 = db.unsqueeze(0).expand(4, nl, nb, N).contiguous()

    def step_mac(terms):
        def f():
            rc = lib.gfhe_ntt_mul_acc(plan.handle, 0, nl, inner, terms, da4.data_ptr(), db4.data_ptr(), nl * inner,
                                      dc.data_ptr(), st)
            if rc:
                raise RuntimeError(_capi.error_string(rc))
        return f

    def step_auto():
        rc = lib.gfhe_automorphism(plan.handle, 0, nl, nb, 5, 1, da.data_ptr(), dc.data_ptr(), st)
        if rc:
            raise RuntimeError(_capi.error_string(rc))

    def step_vec_add():
        rc = lib.gfhe_vec_add(plan.handle, 0, nl, inner, da.data_ptr(), db.data_ptr(), dc.data_ptr(), st)
        if rc:
            raise RuntimeError(_capi.error_string(rc))

    def step_pre():   # poly-mul against a cached NTT(b): db's words stand in for NTT(b) (any canonical residues do)
        rc = lib.gfhe_poly_mul_pre(plan.handle, 0, nl, nb, da.data_ptr(), db.data_ptr(), dc.data_ptr(), st)
        if rc:
            raise RuntimeError(_capi.error_string(rc))

    Ks = max(5, min(K, 30))
    ms_pre = timed(step_pre, Ks, W) / Ks
    ms_mac1 = timed(step_mac(1), Ks, W) / Ks
    ms_mac4 = timed(step_mac(4), Ks, W) / Ks
    ms_auto = timed(step_auto, Ks, W) / Ks
    ms_vadd = timed(step_vec_add, Ks, W) / Ks
    del da4, db4

    # ---- the same shape with 30-bit moduli ("32-bit where the moduli fit"): 64-bit layout and packed int32 layout
    narrow, narrow_err = {}, None
    try:   # an auxiliary section: a failure here is reported, it does not void the headline
        m30 = P.ntt_primes(30, L_TOTAL)[sh.limbs[0]:sh.limbs[1]]
        plan30 = P.make_plan(N, m30, device=local_rank)
        red = torch.tensor(m30, dtype=torch.int64, device=dev).view(nl, 1, 1)
        a30, b30 = da % red, db % red
        c30 = torch.empty_like(a30)
        a30p, b30p = a30.to(torch.int32), b30.to(torch.int32)
        c30p = torch.empty_like(a30p)

        def call30(fn, *ptrs):
            def f():
                rc = fn(plan30.handle, 0, nl, nb, *ptrs)
                if rc:
                    raise RuntimeError(_capi.error_string(rc))
            return f
        for name, fn, ptrs, word in (
                ("poly_mul", lib.gfhe_poly_mul, (a30.data_ptr(), b30.data_ptr(), c30.data_ptr(), ws.data_ptr(), ws_bytes, st), 8),
                ("ntt", lib.gfhe_ntt_forward, (a30.data_ptr(), c30.data_ptr(), st), 8),
                ("intt", lib.gfhe_ntt_inverse, (a30.data_ptr(), c30.data_ptr(), st), 8),
                ("poly_mul_packed", lib.gfhe_poly_mul_u32, (a30p.data_ptr(), b30p.data_ptr(), c30p.data_ptr(), ws.data_ptr(), ws_bytes, st), 4),
                ("ntt_packed", lib.gfhe_ntt_forward_u32, (a30p.data_ptr(), c30p.data_ptr(), st), 4),
                ("intt_packed", lib.gfhe_ntt_inverse_u32, (a30p.data_ptr(), c30p.data_ptr(), st), 4)):
            narrow[name] = (timed(call30(fn, *ptrs), Ks, W) / Ks, word)
        assert torch.equal(c30p.to(torch.int64), c30), "packed and 64-bit layouts disagree"     # both hold iNTT(a30)
        del a30, b30, c30, a30p, b30p, c30p
    except Exception as exc:   # noqa: BLE001
        if world > 1:
            raise          # the timed sections hold barriers: a rank that skipped them would hang the others
        narrow_err = "%s: %s" % (type(exc).__name__, exc)
        log("moduli_30bit section failed:", narrow_err)
    clocks.active = False

    # ---- e2e: host buffers through the C-ABI host entry point (pinned memory, copies inside)
    a_np, b_np, c_np = (t.numpy().view(np.uint64) for t in (pa, pb, pc))
    e2e_steps = max(3, min(K, 20))
    for _ in range(2):
        dispatch.poly_mul_host(a_np, b_np, c_np, moduli, device=local_rank)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        dispatch.poly_mul_host(a_np, b_np, c_np, moduli, device=local_rank)
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    if rank == 0:
        assert np.array_equal(c_np[0, 0], want[0, 0]), "e2e (host-buffer) result is not bit-exact"

    # ---- the one collective: all-gather of result limbs to every rank (reported apart)
    gather_ms = None
    if world > 1:
        for _ in range(2):
            shard.all_gather_limbs(dc, sh)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        full = shard.all_gather_limbs(dc, sh)
        e1.record(stream)
        torch.cuda.synchronize()
        gather_ms = max_over_ranks(e0.elapsed_time(e1))
        assert tuple(full.shape) == (L_TOTAL, B, N)
        del full

    clocks.stop = True
    if prev_affinity:
        os.sched_setaffinity(0, prev_affinity)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    units = UNITS_PER_GPU * world
    value = units * K / (ms_total * 1e-3)
    dom = "k_row_polymul"
    dom_ms = kernels.get(dom, {}).get("ms_per_launch")
    step_kernel_ms = sum(k["ms_per_launch"] * k["launches_per_step"] for k in kernels.values()) or None
    achieved = ALG_BYTES_POLYMUL * UNITS_PER_GPU / (dom_ms * 1e-3) / 1e9 if dom_ms else None
    whole_gbs = ALG_BYTES_POLYMUL * UNITS_PER_GPU / (ms_total / K * 1e-3) / 1e9
    ntt_value = units * K / (ms_ntt * 1e-3)
    ntt_gbs = ALG_BYTES_NTT * UNITS_PER_GPU / (ms_ntt / K * 1e-3) / 1e9

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, per_rep, n_units = cpu_baseline(all_moduli, threads, reps=2)
        cpu = {"value": v, "unit": "limb-poly-mul/s", "cores": threads, "kind": "port",
               "sample": "2 x full config-4 batch [32,16,65536] (%d limb-poly-muls, %.2f s per batch)" % (n_units, per_rep),
               "note": "oracle/ntt_oracle.c (Shoup/Harvey flavour, pthreads over limb-polynomials); the reference "
                       "repo itself has no NTT/poly-mul to time"}

    # the element-wise family the reference DOES implement (arithmetic.py:3-13): its statement `(a + b) % MOD` /
    # `(a * b) % MOD` on dtype=object numpy arrays (the only mode that is exact for 60-bit moduli), one host core, on a
    # 2^20-residue sample of the same operands; next to it the GPU kernels' rate for the same ops
    cpu_vec = None
    if cpu is not None:
        n_s = 1 << 20
        xa = a_h.reshape(-1)[:n_s].astype(object)
        xb = b_h.reshape(-1)[:n_s].astype(object)
        q0 = int(moduli[0])
        t0 = time.perf_counter(); _ = (xa + xb) % q0; t_add = time.perf_counter() - t0
        t0 = time.perf_counter(); _ = (xa * xb) % q0; t_mul = time.perf_counter() - t0
        cpu_vec = {"unit": "residue/s", "cores": 1, "kind": "port",
                   "sample": "2^20 residues of operand a, b (limb 0), numpy dtype=object, statement of arithmetic.py:5,13",
                   "vec_add": n_s / t_add, "vec_mul": n_s / t_mul,
                   "gpu_vec_add": UNITS_PER_GPU * N / (ms_vadd * 1e-3), "gpu_pointwise_mul": UNITS_PER_GPU * N / (ms_mac1 * 1e-3)}

    line = {
        "metric": "limb_poly_mul_per_s", "value": value, "unit": "limb-poly-mul/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_total / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(world),
        "full_rns_poly_mul_per_s": value / L_TOTAL,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak if achieved else None, "traffic": ncu_traffic(dom), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": ALG_BYTES_POLYMUL * UNITS_PER_GPU,
                     "ms_per_launch": dom_ms,
                     "share_of_step": dom_ms / step_kernel_ms if dom_ms and step_kernel_ms else None,
                     "whole_step": {"achieved": whole_gbs, "frac": whole_gbs / peak, "frac_of_8TBs_nominal": whole_gbs / 8000.0}},
        "kernels": kernels,
        "ntt_domain": {
            name: {"ms_per_launch": ms, "bytes_per_residue": bpr, "achieved_GBs": bpr * UNITS_PER_GPU * N / (ms * 1e-3) / 1e9,
                   "frac": bpr * UNITS_PER_GPU * N / (ms * 1e-3) / 1e9 / peak}
            for name, ms, bpr in (("mul_acc_terms1", ms_mac1, 24), ("mul_acc_terms4", ms_mac4, 72),
                                  ("automorphism_ntt", ms_auto, 16), ("vec_add", ms_vadd, 24),
                                  ("poly_mul_pre (one operand cached in NTT form)", ms_pre, 24))},
        "moduli_30bit": {
            name: {"ms_per_launch": ms, "limb_ops_per_s": UNITS_PER_GPU * world / (ms * 1e-3),
                   "bytes_per_residue": (3 if name.startswith("poly_mul") else 2) * word,
                   "achieved_GBs": (3 if name.startswith("poly_mul") else 2) * word * N * UNITS_PER_GPU / (ms * 1e-3) / 1e9,
                   "frac": (3 if name.startswith("poly_mul") else 2) * word * N * UNITS_PER_GPU / (ms * 1e-3) / 1e9 / peak}
            for name, (ms, word) in narrow.items()} if narrow_err is None else {"error": narrow_err},
        "ntt": {"metric": "limb_ntt_per_s", "value": ntt_value, "unit": "limb-NTT/s", "ms_per_step": ms_ntt / K,
                "achieved_GBs": ntt_gbs, "frac": ntt_gbs / peak, "frac_of_8TBs_nominal": ntt_gbs / 8000.0,
                "inverse": {"value": units * K / (ms_intt * 1e-3), "ms_per_step": ms_intt / K,
                            "frac": ALG_BYTES_NTT * UNITS_PER_GPU / (ms_intt / K * 1e-3) / 1e9 / peak}},
        "e2e": {"value": units * e2e_steps / e2e_s, "unit": "limb-poly-mul/s", "steps": e2e_steps,
                "ms_per_step": e2e_s / e2e_steps * 1e3,
                "h2d_bytes_per_step": int(2 * a_h.nbytes), "d2h_bytes_per_step": int(a_h.nbytes),
                "api": "gfhe_poly_mul_host (pinned host arrays; limb groups pipelined H2D/kernels/D2H)"},
        "cpu_baseline_vec_ops": cpu_vec,
        "gather": {"all_gather_ms": gather_ms, "bytes_out_per_rank": int(a_h.nbytes) * world} if world > 1 else None,
        "gpu_launches": int(round(sum(k["launches_per_step"] for k in kernels.values()) * K)) if kernels else 3 * K,
        "clocks": clocks.summary(),
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        log("bench.py: --gpus %d needs torchrun (one rank per GPU); running 1 rank" % args.gpus)
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
