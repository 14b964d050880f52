#!/usr/bin/env python
"""Benchmark of the NTT / poly-mul hot path (BASELINE.json metric: NTT/s and poly-mul/s at N=2^16).

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload = BASELINE.json configs[3] AS STATED ("config 4"): negacyclic poly-mul at N=2^16 over 32 RNS limbs of 60-bit
primes, batch 16 -- a FIXED global [32, 16, 65536] array (512 limb-poly-muls, 805.3 MB algorithmic), partitioned by
shard.partition(32, 16, n_gpus, rank) (limbs first: 32/n limbs x 16 batch per GPU), so the scaling is STRONG: the
whole job is the same 512 units at every GPU count.  Compute needs no collective; the one collective is the final
all-gather of result limbs to every rank, reported both apart (`gather`) and inside (`value_with_gather`).

One JSON line on stdout (rank 0).  value = global limb-poly-muls/s with operands resident in HBM, timed with CUDA
events on the launching stream, max over ranks.  e2e = the same through the host-buffer C-ABI call (pinned host arrays,
H2D + kernels + D2H inside the timed region).  extra.configs holds BASELINE.json's other configs (2, 3, 5) measured the
same way; extra.weak the per-GPU-work-fixed variant round 1 reported.  Every multi-GPU result is gathered and compared
with the oracle on every rank before anything is timed.
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
B_TOTAL = 16
BITS = 60
UNITS_TOTAL = L_TOTAL * B_TOTAL               # limb-poly-muls in the whole job (every GPU count)
ALG_BYTES_POLYMUL = 24 * N                     # per limb-poly-mul: read a, read b, write c
ALG_BYTES_NTT = 16 * N                         # per limb-NTT
FALLBACK_HBM_GBS = 6650.0                      # B200_PROFILING.md fallback
L2_BYTES = 126e6
SEED = 20261104 + 4                            # seed = 20261104 + config index


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# stdout carries exactly ONE JSON line.  Libraries write there too (NCCL prints its version line to stdout), so file
# descriptor 1 is pointed at stderr for the life of the process and the line goes to a private copy of the real stdout.
_REAL_STDOUT = None


def claim_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """dram bytes per launch of `kernel` (full config-4 launch on one GPU) from the committed ncu summary, or None."""
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


def limb_block(limb, q, which, B=B_TOTAL, n=N, seed=SEED):
    """Operand `which` (0 = a, 1 = b) of one limb, full batch: every rank can regenerate any limb of the global input."""
    import numpy as np
    return np.random.default_rng([seed, which, limb]).integers(0, q, (B, n), dtype=np.uint64)


def shard_inputs(moduli, sh, B=B_TOTAL, n=N, seed=SEED):
    import numpy as np
    a = np.stack([limb_block(l, moduli[l], 0, B, n, seed)[sh.batch[0]:sh.batch[1]] for l in range(*sh.limbs)])
    b = np.stack([limb_block(l, moduli[l], 1, B, n, seed)[sh.batch[0]:sh.batch[1]] for l in range(*sh.limbs)])
    return a, b


def cpu_polymul_rate(moduli, threads, reps, warm=1):
    """The CPU arm of both legs (cpu_baseline of the B200 line, and --impl reference): the C oracle port
    (oracle/ntt_oracle.c -- the reference itself has no NTT), Shoup/Harvey flavour, pthreads over limb-polynomials,
    full config-4 batches: `warm` untimed batches (page-in, thread spin-up), then `reps` timed ones."""
    import numpy as np
    from oracle import cref
    from oracle import oracle as O
    cref.build()
    R = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    a = np.stack([limb_block(l, moduli[l], 0) for l in range(len(moduli))])
    b = np.stack([limb_block(l, moduli[l], 1) for l in range(len(moduli))])
    for _ in range(max(warm, 1)):
        R.poly_mul(a, b, fast=True, nthreads=threads)
    t0 = time.perf_counter()
    for _ in range(reps):
        c = R.poly_mul(a, b, fast=True, nthreads=threads)
    dt = time.perf_counter() - t0
    assert c.shape == a.shape and np.all(c[0, 0] < moduli[0])
    return UNITS_TOTAL * reps / dt, dt / reps


def _shard_module():
    """gpu-fhe_b200/shard.py without the package around it: the reference arm must not load the CUDA library."""
    if "gpu_fhe_b200" in sys.modules:
        from gpu_fhe_b200 import shard
        return shard
    import importlib.util
    spec = importlib.util.spec_from_file_location("_gfhe_shard_only", os.path.join(ROOT, "gpu-fhe_b200", "shard.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod          # dataclasses looks the module up by name
    spec.loader.exec_module(mod)
    return mod


def workload_config(n_gpus):
    shard = _shard_module()
    sh = shard.partition(L_TOTAL, B_TOTAL, max(n_gpus, 1), 0)
    set_bytes = 3 * sh.num_limbs * sh.num_batch * N * 8
    nset = buffer_sets(set_bytes)
    return {"workload": "config 4: negacyclic poly-mul, global [32,16,65536] (32 limbs of 60-bit primes, batch 16, N=65536), "
                        "%d limbs x %d batch per GPU on %d GPU%s" % (sh.num_limbs, sh.num_batch, n_gpus, "" if n_gpus == 1 else "s"),
            "N": N, "limbs": L_TOTAL, "batch": B_TOTAL, "modulus_bits": BITS,
            "units_per_step": UNITS_TOTAL, "units_per_gpu_per_step": sh.num_limbs * sh.num_batch,
            "parallelism": "limb-shard x%d%s" % (sh.g_l, " batch-shard x%d" % sh.g_b if sh.g_b > 1 else ""),
            "l2_policy": ("inputs larger than L2 (%.0f MB touched per step per GPU vs 126 MB L2); no flush" % (set_bytes * 4 / 3 / 1e6)) if nset == 1 else
                         ("per-GPU shard (%.0f MB per step) is not far above the 126 MB L2: steps rotate over %d distinct buffer sets "
                          "(%.0f MB) so no step finds its operands in L2" % (set_bytes * 4 / 3 / 1e6, nset, nset * set_bytes * 4 / 3 / 1e6))}


def buffer_sets(set_bytes):
    """How many distinct (a, b, c) sets the timed loop rotates over so that consecutive steps never hit L2."""
    if set_bytes >= 3 * L2_BYTES:
        return 1
    return int(min(8, max(2, -(-3 * L2_BYTES // set_bytes))))


def run_reference(args, rank, world):
    """Reference arm: the path's CPU implementation on the box's host cores.  The reference repo has no NTT / poly-mul
    (identity stubs, /root/reference/arithmetic.py:15-19), so this is the oracle port (kind = "port").  Rank 0 only."""
    if rank != 0:
        return
    from oracle import oracle as O
    threads = os.cpu_count() or 1
    moduli = O.ntt_primes(BITS, L_TOTAL)
    value, per_step = cpu_polymul_rate(moduli, threads, reps=args.steps, warm=max(args.warmup, 1))
    sample = "each step = one full config-4 batch [32,16,65536] (512 limb-poly-muls) on %d host threads; %d untimed warm-up batches" % (
        threads, max(args.warmup, 1))
    line = {
        "impl": "reference", "metric": "limb_poly_mul_per_s", "value": value, "unit": "limb-poly-mul/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": value, "unit": "limb-poly-mul/s", "cores": threads, "kind": "port", "sample": sample,
                         "note": "reference repo has no NTT/poly-mul (arithmetic.py:15-19 are identity stubs); "
                                 "this is oracle/ntt_oracle.c, Shoup/Harvey flavour, pthreads over limb-polynomials"},
        "e2e": {"value": value, "unit": "limb-poly-mul/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


class Ctx:
    """Per-rank plumbing: device, stream, barriers, max-over-ranks, event timing of a step function."""

    def __init__(self, rank, world, local_rank):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        from gpu_fhe_b200 import shard
        # the block of GPUs with the fastest host links when the box shows at least twice as many GPUs as the job has ranks
        # (measured: 4 ranks on GPUs 4-7 of this pool's 8-GPU boxes move host data 1.9x faster than on 0-3, e2e 7.13 ->
        # 4.21 ms, profiles/r02_run23_*); GFHE_DEVICE_MAP=identity keeps device = local rank
        os.environ.setdefault("GFHE_DEVICE_MAP", "probe")
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world))
        local_rank, self.device_probe = shard.choose_local_devices(local_world, local_rank)
        self.rank, self.world, self.local = rank, world, local_rank
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        if world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.stream = torch.cuda.current_stream(self.dev)
        import ctypes
        self.st = ctypes.c_void_p(self.stream.cuda_stream)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def all_true(self, ok):
        return self.max_over_ranks(0.0 if ok else 1.0) == 0.0

    def timed(self, step, steps, warmup, active=True):
        """ms for `steps` calls of step(i) after `warmup` untimed ones: CUDA events on the launching stream, barrier +
        synchronize on both sides, max over ranks.  Ranks with active=False do no work but keep the collectives."""
        torch = self.torch
        if active:
            for i in range(warmup):
                step(i)
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        if active:
            for i in range(steps):
                step(warmup + i)
        e1.record(self.stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) if active else 0.0
        self.barrier()
        return self.max_over_ranks(ms)


def check(rc, what="gfhe call"):
    if rc:
        from gpu_fhe_b200 import _capi
        raise RuntimeError("%s: %s" % (what, _capi.error_string(rc)))


def subgroup_size(world, L, B):
    """GPUs a [L, B, N] job can use without splitting a transform: the largest power of two <= min(world, L*B)
    that partitions it evenly."""
    from gpu_fhe_b200 import shard
    g = 1
    while g * 2 <= min(world, L * B):
        g *= 2
    while g > 1:
        try:
            shard.partition(L, B, g, 0)
            return g
        except ValueError:
            g //= 2
    return 1


def run_config(cx, lib, P, name, op, n, L, B, bits, steps, warmup, peak, gpus=None):
    """One of BASELINE.json's other configs, device-resident, on `gpus` ranks (default: as many as the job can use):
    [L, B, n] sharded limbs-first, buffers rotated past L2, CUDA events, max over ranks.  Inputs are uniform in [0, 2^(bits-1))
    (below every modulus of the family), generated on the device."""
    import torch
    from gpu_fhe_b200 import shard
    g = gpus or subgroup_size(cx.world, L, B)
    active = cx.rank < g
    res = {"name": name, "op": op, "N": n, "L": L, "B": B, "bits": bits, "gpus": g}
    ms = None
    if active:
        sh = shard.partition(L, B, g, cx.rank)
        nl, nb = sh.num_limbs, sh.num_batch
        moduli = P.ntt_primes(bits, L)[sh.limbs[0]:sh.limbs[1]]
        plan = P.make_plan(n, moduli, device=cx.local)
        words = nl * nb * n
        nbuf = 3 if op == "polymul" else 2
        nset = buffer_sets(nbuf * words * 8)
        gen = torch.Generator(device=cx.dev)
        gen.manual_seed(20261104 + n + 131 * L + B)
        a = [torch.randint(0, 1 << (bits - 1), (nl, nb, n), dtype=torch.int64, device=cx.dev, generator=gen) for _ in range(nset)]
        b = [torch.randint(0, 1 << (bits - 1), (nl, nb, n), dtype=torch.int64, device=cx.dev, generator=gen) for _ in range(nset)] if op == "polymul" else None
        c = [torch.empty_like(a[0]) for _ in range(nset)]
        ws_bytes = lib.gfhe_poly_mul_workspace_bytes(plan.handle, nl, nb) if op == "polymul" else 0
        ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=cx.dev)

        def step(i):
            k = i % nset
            if op == "polymul":
                check(lib.gfhe_poly_mul(plan.handle, 0, nl, nb, a[k].data_ptr(), b[k].data_ptr(), c[k].data_ptr(), ws.data_ptr(), ws_bytes, cx.st), name)
            elif op == "ntt":
                check(lib.gfhe_ntt_forward(plan.handle, 0, nl, nb, a[k].data_ptr(), c[k].data_ptr(), cx.st), name)
            else:
                check(lib.gfhe_ntt_inverse(plan.handle, 0, nl, nb, a[k].data_ptr(), c[k].data_ptr(), cx.st), name)
        res["buffer_sets"] = nset
    else:
        def step(i):
            pass
    ms = cx.timed(step, steps, warmup, active) / steps
    if active:
        del a, b, c, ws
    units = L * B
    alg = (24 if op == "polymul" else 16) * n * units
    res.update({"ms_per_step": ms, "limb_ops_per_s": units / (ms * 1e-3), "algorithmic_GBs": alg / (ms * 1e-3) / 1e9,
                "frac_of_aggregate_hbm": alg / (ms * 1e-3) / 1e9 / (g * peak)})
    return res


def run_b200(args, rank, world, local_rank):
    import ctypes

    import numpy as np
    import torch

    import polynomial as P
    from gpu_fhe_b200 import _capi, dispatch, shard

    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    cx = Ctx(rank, world, local_rank)
    dev, st, dist = cx.dev, cx.st, cx.dist
    local_rank = cx.local
    lib = _capi.lib
    peak, peak_src = measured_peak()

    # several ranks on one box: keep each rank's pinned buffers and copy threads on its GPU's NUMA node (e2e leg)
    prev_affinity = shard.bind_to_gpu_numa(local_rank) if world > 1 and not os.environ.get("GFHE_BENCH_NO_NUMA") else None
    sh = shard.partition(L_TOTAL, B_TOTAL, world, rank)
    all_moduli = P.ntt_primes(BITS, L_TOTAL)
    moduli = all_moduli[sh.limbs[0]:sh.limbs[1]]
    nl, nb = sh.num_limbs, sh.num_batch
    units_local = nl * nb
    plan = P.make_plan(N, moduli, device=local_rank)

    a_h, b_h = shard_inputs(all_moduli, sh)
    pa = torch.from_numpy(a_h.view(np.int64)).pin_memory()
    pb = torch.from_numpy(b_h.view(np.int64)).pin_memory()
    pc = torch.empty_like(pa).pin_memory()
    nset = buffer_sets(3 * a_h.nbytes)
    da = [pa.to(dev) for _ in range(nset)]
    db = [pb.to(dev) for _ in range(nset)]
    dc = [torch.empty_like(da[0]) for _ in range(nset)]
    ws_bytes = lib.gfhe_poly_mul_workspace_bytes(plan.handle, nl, nb)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)

    def step_polymul(i=0):
        k = i % nset
        check(lib.gfhe_poly_mul(plan.handle, 0, nl, nb, da[k].data_ptr(), db[k].data_ptr(), dc[k].data_ptr(), ws.data_ptr(), ws_bytes, st), "gfhe_poly_mul")

    def step_ntt(i=0):
        k = i % nset
        check(lib.gfhe_ntt_forward(plan.handle, 0, nl, nb, da[k].data_ptr(), dc[k].data_ptr(), st), "gfhe_ntt_forward")

    def step_intt(i=0):
        k = i % nset
        check(lib.gfhe_ntt_inverse(plan.handle, 0, nl, nb, da[k].data_ptr(), dc[k].data_ptr(), st), "gfhe_ntt_inverse")

    clocks = ClockSampler(local_rank)
    K, W = args.steps, max(args.warmup, 3)

    # ---- parity before anything is timed: gather the N-GPU result and compare it with the oracle (checker only) ----------
    # every rank: one limb-polynomial out of EVERY rank's shard;  rank 0: the whole [32,16,65536] array
    from oracle import cref
    from oracle import oracle as O
    step_polymul(0)
    full = shard.all_gather_limbs(dc[0], sh) if world > 1 else dc[0]
    torch.cuda.synchronize()
    assert tuple(full.shape) == (L_TOTAL, B_TOTAL, N)
    ok = True
    for r in range(world):
        o = shard.partition(L_TOTAL, B_TOTAL, world, r)
        l, j = o.limbs[0] + (rank + r) % o.num_limbs, o.batch[0] + (3 * rank + r) % o.num_batch
        R1 = cref.Ring(N, [all_moduli[l]], [O.find_psi(all_moduli[l], N)])
        want = R1.poly_mul(limb_block(l, all_moduli[l], 0)[None, j:j + 1], limb_block(l, all_moduli[l], 1)[None, j:j + 1], fast=True)[0, 0]
        ok = ok and np.array_equal(full[l, j].cpu().numpy().view(np.uint64), want)
    parity = {"every_rank_checks_one_unit_of_every_shard": None, "rank0_full_array": None}
    if rank == 0:
        Rf = cref.Ring(N, all_moduli, [O.find_psi(q, N) for q in all_moduli])
        fa = np.stack([limb_block(l, all_moduli[l], 0) for l in range(L_TOTAL)])
        fb = np.stack([limb_block(l, all_moduli[l], 1) for l in range(L_TOTAL)])
        want_full = Rf.poly_mul(fa, fb, fast=True, nthreads=os.cpu_count() or 1)
        full_ok = np.array_equal(full.cpu().numpy().view(np.uint64), want_full)
        parity["rank0_full_array"] = bool(full_ok)
        ok = ok and full_ok
        want_local = want_full[sh.limbs[0]:sh.limbs[1], sh.batch[0]:sh.batch[1]].copy()
        del fa, fb
        if world == 1:
            del want_full
    parity["every_rank_checks_one_unit_of_every_shard"] = cx.all_true(ok)
    assert parity["every_rank_checks_one_unit_of_every_shard"], "gathered %d-GPU result is not bit-exact against the oracle" % world
    del full

    # ---- headline: device-resident poly-mul of the fixed global job (strong scaling)
    clocks.active = True
    ms_total = cx.timed(step_polymul, K, W)
    # ---- per-kernel live durations (events around every launch, same loop)
    lib.gfhe_profile_enable(plan.handle, 1)
    cx.timed(step_polymul, K, 1)
    names = (ctypes.c_char_p * 8)()
    tot = (ctypes.c_double * 8)()
    cnt = (ctypes.c_uint64 * 8)()
    nk = lib.gfhe_profile_read(plan.handle, 8, names, tot, cnt)
    lib.gfhe_profile_enable(plan.handle, 0)
    kernels = {names[i].decode(): {"ms_per_launch": tot[i] / cnt[i], "launches_per_step": cnt[i] / (K + 1)} for i in range(max(nk, 0))}
    # ---- forward / inverse NTT throughput (second half of the metric), same shard
    ms_ntt = cx.timed(step_ntt, K, W)
    ms_intt = cx.timed(step_intt, K, W)

    # ---- the one collective: all-gather of result limbs to every rank -- apart, and inside the step
    gather = None
    if world > 1:
        Kg = max(5, min(K, 20))

        def step_gather(i=0):
            shard.all_gather_limbs(dc[i % nset], sh)

        def step_both(i=0):
            step_polymul(i)
            shard.all_gather_limbs(dc[i % nset], sh)
        ms_gather = cx.timed(step_gather, Kg, 3) / Kg
        ms_both = cx.timed(step_both, Kg, 3) / Kg
        out_bytes = L_TOTAL * B_TOTAL * N * 8
        gather = {"all_gather_ms": ms_gather, "step_with_gather_ms": ms_both, "bytes_out_per_rank": out_bytes,
                  "bytes_received_per_rank": out_bytes - int(a_h.nbytes),
                  "inbound_GBs_per_rank": (out_bytes - a_h.nbytes) / (ms_gather * 1e-3) / 1e9,
                  "note": "torch.distributed.all_gather_into_tensor over NCCL after the kernels, same stream order (no overlap)"}

    # ---- the same step with the gather OVERLAPPED: cyclic limb map (rank r holds limbs r, r+n, ...), so the k-th local limbs
    # of all ranks are a contiguous slab of the result and their all-gather can run while limb k+1 is being multiplied
    if world > 1:
        csh = shard.partition(L_TOTAL, B_TOTAL, world, rank, cyclic=True)
        cmod = [all_moduli[l] for l in csh.limb_ids]
        cplan = P.make_plan(N, cmod, device=local_rank)
        ca = np.stack([limb_block(l, all_moduli[l], 0)[csh.batch[0]:csh.batch[1]] for l in csh.limb_ids])
        cb = np.stack([limb_block(l, all_moduli[l], 1)[csh.batch[0]:csh.batch[1]] for l in csh.limb_ids])
        dca, dcb = torch.from_numpy(ca.view(np.int64)).to(dev), torch.from_numpy(cb.view(np.int64)).to(dev)
        dcc = torch.empty_like(dca)
        full_c = torch.empty((L_TOTAL, B_TOTAL, N), dtype=torch.int64, device=dev)
        cnl, cnb = csh.num_limbs, csh.num_batch
        limb_bytes = cnb * N * 8
        ws1 = lib.gfhe_poly_mul_workspace_bytes(cplan.handle, 1, cnb)

        def step_overlap(i=0):
            works = []
            for k in range(cnl):
                o = k * limb_bytes
                check(lib.gfhe_poly_mul(cplan.handle, k, 1, cnb, dca.data_ptr() + o, dcb.data_ptr() + o, dcc.data_ptr() + o, ws.data_ptr(), ws1, st), "overlap")
                works.append(shard.all_gather_limb_async(dcc[k], full_c, k, csh))
            for w in works:
                w.wait()
        step_overlap()
        torch.cuda.synchronize()
        ok = True
        got = full_c.cpu().numpy().view(np.uint64)
        if rank == 0:
            ok = np.array_equal(got, want_full)
        else:                                          # one unit of every limb slab, against the oracle
            for k in range(cnl):
                l, j = k * csh.g_l + (rank + k) % csh.g_l, (5 * rank + k) % B_TOTAL
                R1 = cref.Ring(N, [all_moduli[l]], [O.find_psi(all_moduli[l], N)])
                want = R1.poly_mul(limb_block(l, all_moduli[l], 0)[None, j:j + 1], limb_block(l, all_moduli[l], 1)[None, j:j + 1], fast=True)[0, 0]
                ok = ok and np.array_equal(got[l, j], want)
        parity["overlapped_gather_result"] = cx.all_true(ok)
        assert parity["overlapped_gather_result"], "overlapped per-limb gather is not bit-exact against the oracle"
        ms_overlap = cx.timed(step_overlap, Kg, 3) / Kg
        gather["step_with_overlapped_gather_ms"] = ms_overlap
        gather["overlap_note"] = ("cyclic limb map: poly-mul of local limb k, then its all-gather (async, NCCL's stream) into the contiguous "
                                  "slab [k*n, (k+1)*n) of the result while limb k+1 is multiplied")
        del dca, dcb, dcc, full_c, got

    # ---- the gather as NVLink peer stores issued by the library itself (csrc/gather.cu, shard.PeerGather): each rank pushes its
    # block into every rank's full array; one stream-ordered flag all-reduce closes the step.  Whole block after the kernels,
    # and chunk by chunk on a side stream while the next chunk is multiplied.
    pgx, peer_err = None, None
    if world > 1:
        try:                                          # CUDA IPC between the ranks' processes: report, do not die, if a box forbids it
            pgx = shard.PeerGather(L_TOTAL, B_TOTAL, N, dev)
        except Exception as exc:   # noqa: BLE001
            peer_err = "%s: %s" % (type(exc).__name__, exc)
            log("peer-store gather unavailable on rank %d: %s" % (rank, peer_err))
        if not cx.all_true(pgx is not None):          # all ranks or none (close() is collective: a rank that got one just drops it)
            pgx = None
            gather["peer_store"] = {"error": peer_err or "unavailable on another rank"}
    if pgx is not None:
        side = torch.cuda.Stream(dev)
        limb_bytes = nb * N * 8

        mine = pgx.local_block(sh)                    # the kernels write this rank's block straight into its own full array

        def step_push_only(i=0):
            pgx.push(mine, sh)
            pgx.finish()

        def step_push(i=0):
            k = i % nset
            check(lib.gfhe_poly_mul(plan.handle, 0, nl, nb, da[k].data_ptr(), db[k].data_ptr(), mine.data_ptr(), ws.data_ptr(), ws_bytes, st), "gfhe_poly_mul")
            pgx.push(mine, sh)
            pgx.finish()

        def make_overlap(cl):
            def f(i=0):
                ks = i % nset
                wsc = lib.gfhe_poly_mul_workspace_bytes(plan.handle, cl, nb)
                for k0 in range(0, nl, cl):
                    o = k0 * limb_bytes
                    check(lib.gfhe_poly_mul(plan.handle, k0, cl, nb, da[ks].data_ptr() + o, db[ks].data_ptr() + o, mine.data_ptr() + o,
                                            ws.data_ptr(), wsc, st), "push overlap")
                    ev = torch.cuda.Event()
                    ev.record(cx.stream)
                    side.wait_event(ev)
                    with torch.cuda.stream(side):
                        for k in range(k0, k0 + cl):
                            pgx.push_limb(mine, k, sh)
                ev = torch.cuda.Event()
                ev.record(side)
                cx.stream.wait_event(ev)
                pgx.finish()
            return f
        pgx.full.zero_()
        cx.barrier()
        step_push(0)
        torch.cuda.synchronize()
        got = pgx.full.cpu().numpy().view(np.uint64)
        ok = True
        if rank == 0:
            ok = np.array_equal(got, want_full)
        else:
            for r in range(world):
                o = shard.partition(L_TOTAL, B_TOTAL, world, r)
                l, j = o.limbs[0] + (2 * rank + r) % o.num_limbs, o.batch[0] + (7 * rank + r) % o.num_batch
                R1 = cref.Ring(N, [all_moduli[l]], [O.find_psi(all_moduli[l], N)])
                want = R1.poly_mul(limb_block(l, all_moduli[l], 0)[None, j:j + 1], limb_block(l, all_moduli[l], 1)[None, j:j + 1], fast=True)[0, 0]
                ok = ok and np.array_equal(got[l, j], want)
        parity["peer_store_gather_result"] = cx.all_true(ok)
        assert parity["peer_store_gather_result"], "peer-store gather is not bit-exact against the oracle"
        ref_full = pgx.full.clone()                   # the checked result: the overlapped variants must reproduce it
        ms_push_only = cx.timed(step_push_only, Kg, 3) / Kg
        ms_push = cx.timed(step_push, Kg, 3) / Kg
        best_cl, ms_push_ov = None, None
        for cl in (1,):
            f = make_overlap(cl)
            pgx.full.zero_()
            cx.barrier()
            f(0)
            torch.cuda.synchronize()
            same = rank != 0 or np.array_equal(pgx.full.cpu().numpy().view(np.uint64), want_full)
            if rank != 0:      # every other rank: its own block and one limb of a neighbour's, against what the kernels wrote before
                nb_ = shard.partition(L_TOTAL, B_TOTAL, world, (rank + 1) % world)
                same = bool(torch.equal(pgx.full[sh.limbs[0]:sh.limbs[1]], dc[0])) and bool(torch.equal(pgx.full[nb_.limbs[0]], ref_full[nb_.limbs[0]]))
            assert cx.all_true(same), "overlapped peer-store gather differs"
            ms = cx.timed(f, Kg, 3) / Kg
            if ms_push_ov is None or ms < ms_push_ov:
                best_cl, ms_push_ov = cl, ms
        gather["peer_store"] = {"gather_ms": ms_push_only, "step_with_gather_ms": ms_push, "step_with_overlapped_gather_ms": ms_push_ov,
                                "overlap_chunk_limbs": best_cl,
                                "inbound_GBs_per_rank": (L_TOTAL * B_TOTAL * N * 8 - a_h.nbytes) / (ms_push_only * 1e-3) / 1e9,
                                "note": "gfhe_push_gather: 16-byte stores on CUDA-IPC peer pointers over NVLink, this rank's block into every rank's "
                                        "[32,16,65536] array, then one flag all-reduce; overlapped = chunk k pushed on a side stream under chunk k+1's kernels"}
        del got, ref_full, mine
        pgx.close()
    if world > 1 and rank == 0:
        del want_full

    # ---- round 1's weak-scaling variant (per-GPU work fixed at 512 units: 32/n limbs x 16n batch per rank)
    weak = None
    if world > 1:
        wsh = shard.partition(L_TOTAL, B_TOTAL * world, world, rank)
        wl, wb = wsh.num_limbs, wsh.num_batch
        gen = torch.Generator(device=dev)
        gen.manual_seed(7 + rank)
        wa = torch.randint(0, 1 << (BITS - 1), (wl, wb, N), dtype=torch.int64, device=dev, generator=gen)
        wb_ = torch.randint(0, 1 << (BITS - 1), (wl, wb, N), dtype=torch.int64, device=dev, generator=gen)
        wc = torch.empty_like(wa)
        wws_bytes = lib.gfhe_poly_mul_workspace_bytes(plan.handle, wl, wb)
        wws = torch.empty(wws_bytes, dtype=torch.uint8, device=dev)

        def step_weak(i=0):
            check(lib.gfhe_poly_mul(plan.handle, 0, wl, wb, wa.data_ptr(), wb_.data_ptr(), wc.data_ptr(), wws.data_ptr(), wws_bytes, st), "weak")
        Kw = max(5, min(K, 30))
        ms_weak = cx.timed(step_weak, Kw, W) / Kw
        weak = {"scaling": "weak", "workload": "global [32, %d, 65536]: 512 limb-poly-muls per GPU" % (B_TOTAL * world),
                "value": UNITS_TOTAL * world / (ms_weak * 1e-3), "unit": "limb-poly-mul/s", "ms_per_step": ms_weak}
        del wa, wb_, wc, wws

    # ---- auxiliary sections at one GPU: NTT-domain-resident kernels and the 30-bit layouts
    ntt_domain, narrow, narrow_err = None, {}, None
    Ks = max(5, min(K, 30))
    if world == 1:
        inner = nb * N
        da4 = da[0].unsqueeze(0).expand(4, nl, nb, N).contiguous()
        db4 = db[0].unsqueeze(0).expand(4, nl, nb, N).contiguous()

        def step_mac(terms):
            def f(i=0):
                check(lib.gfhe_ntt_mul_acc(plan.handle, 0, nl, inner, terms, da4.data_ptr(), db4.data_ptr(), nl * inner, dc[0].data_ptr(), st))
            return f

        def step_auto(i=0):
            check(lib.gfhe_automorphism(plan.handle, 0, nl, nb, 5, 1, da[0].data_ptr(), dc[0].data_ptr(), st))

        def step_vec_add(i=0):
            check(lib.gfhe_vec_add(plan.handle, 0, nl, inner, da[0].data_ptr(), db[0].data_ptr(), dc[0].data_ptr(), st))

        def step_pre(i=0):   # poly-mul against a cached NTT(b): db's words stand in for NTT(b) (any canonical residues do)
            check(lib.gfhe_poly_mul_pre(plan.handle, 0, nl, nb, da[0].data_ptr(), db[0].data_ptr(), dc[0].data_ptr(), st))
        rows = (("mul_acc_terms1", step_mac(1), 24), ("mul_acc_terms4", step_mac(4), 72), ("automorphism_ntt", step_auto, 16),
                ("vec_add", step_vec_add, 24), ("poly_mul_pre (one operand cached in NTT form)", step_pre, 24))
        ntt_domain = {}
        for name, fn, bpr in rows:
            ms = cx.timed(fn, Ks, W) / Ks
            gbs = bpr * units_local * N / (ms * 1e-3) / 1e9
            ntt_domain[name] = {"ms_per_launch": ms, "bytes_per_residue": bpr, "achieved_GBs": gbs, "frac": gbs / peak}
        del da4, db4
        try:   # the same shape with 30-bit moduli ("32-bit where the moduli fit"): 64-bit layout and packed int32 layout
            m30 = P.ntt_primes(30, L_TOTAL)
            plan30 = P.make_plan(N, m30, device=local_rank)
            red = torch.tensor(m30, dtype=torch.int64, device=dev).view(nl, 1, 1)
            a30, b30 = da[0] % red, db[0] % red
            c30 = torch.empty_like(a30)
            a30p, b30p = a30.to(torch.int32), b30.to(torch.int32)
            c30p = torch.empty_like(a30p)

            def call30(fn, *ptrs):
                def f(i=0):
                    check(fn(plan30.handle, 0, nl, nb, *ptrs))
                return f
            for name, fn, ptrs, word in (
                    ("poly_mul", lib.gfhe_poly_mul, (a30.data_ptr(), b30.data_ptr(), c30.data_ptr(), ws.data_ptr(), ws_bytes, st), 8),
                    ("ntt", lib.gfhe_ntt_forward, (a30.data_ptr(), c30.data_ptr(), st), 8),
                    ("intt", lib.gfhe_ntt_inverse, (a30.data_ptr(), c30.data_ptr(), st), 8),
                    ("poly_mul_packed", lib.gfhe_poly_mul_u32, (a30p.data_ptr(), b30p.data_ptr(), c30p.data_ptr(), ws.data_ptr(), ws_bytes, st), 4),
                    ("ntt_packed", lib.gfhe_ntt_forward_u32, (a30p.data_ptr(), c30p.data_ptr(), st), 4),
                    ("intt_packed", lib.gfhe_ntt_inverse_u32, (a30p.data_ptr(), c30p.data_ptr(), st), 4)):
                ms = cx.timed(call30(fn, *ptrs), Ks, W) / Ks
                bpr = (3 if name.startswith("poly_mul") else 2) * word
                gbs = bpr * N * units_local / (ms * 1e-3) / 1e9
                narrow[name] = {"ms_per_launch": ms, "limb_ops_per_s": units_local / (ms * 1e-3), "bytes_per_residue": bpr,
                                "achieved_GBs": gbs, "frac": gbs / peak}
            assert torch.equal(c30p.to(torch.int64), c30), "packed and 64-bit layouts disagree"     # both hold iNTT(a30)
            del a30, b30, c30, a30p, b30p, c30p
        except Exception as exc:   # noqa: BLE001 -- an auxiliary section: a failure is reported, it does not void the headline
            narrow_err = "%s: %s" % (type(exc).__name__, exc)
            log("moduli_30bit section failed:", narrow_err)

    # ---- BASELINE.json's other configs, same method (device-resident, events, buffers rotated past L2)
    del da[1:], db[1:], dc[1:]
    torch.cuda.empty_cache()
    configs = []
    Kc, Wc = max(5, min(K, 20)), 3
    if world == 1:
        configs.append(run_config(cx, lib, P, "config 2: N=4096 poly-mul, 4 limbs of 50-bit primes, batch 256, 1 GPU", "polymul", 4096, 4, 256, 50, Kc, Wc, peak, gpus=1))
    if world <= 2:
        for op in ("ntt", "intt"):
            configs.append(run_config(cx, lib, P, "config 3: N=16384 batched %s, 8 limbs of 60-bit primes, batch 128, %d GPU%s (limb-sharded)"
                                      % (op.upper(), world, "" if world == 1 else "s"), op, 16384, 8, 128, 60, Kc, Wc, peak, gpus=world))
    sweep = []
    if not args.no_sweep:
        for Ls in (1, 2, 4, 8, 16, 32, 64):
            for Bs in (1, 2, 4, 8, 16, 32, 64):
                r = run_config(cx, lib, P, "config 5", "ntt", 1 << 17, Ls, Bs, 60, max(5, min(K, 10)), 3, peak)
                sweep.append({k: r[k] for k in ("L", "B", "gpus", "ms_per_step", "limb_ops_per_s", "frac_of_aggregate_hbm")})
        P.clear_plans()
        plan = P.make_plan(N, moduli, device=local_rank)
    clocks.active = False

    # ---- e2e: host buffers through the C-ABI host entry point (copies inside the timed region)
    a_np, b_np, c_np = (t.numpy().view(np.uint64) for t in (pa, pb, pc))
    e2e_steps = max(3, min(K, 20))

    own = {}

    def e2e_leg(a, b, c, steps, mod=None):
        mod = moduli if mod is None else mod
        for _ in range(2):
            dispatch.poly_mul_host(a, b, c, mod, device=local_rank)
        cx.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            dispatch.poly_mul_host(a, b, c, mod, device=local_rank)
        own["s"] = (time.perf_counter() - t0) / steps          # this rank's own time per call
        s = cx.max_over_ranks(own["s"] * steps)
        hs = dispatch.host_stats(mod, N, device=local_rank)
        cx.barrier()
        return s / steps, hs
    e2e_s, hs = e2e_leg(a_np, b_np, c_np, e2e_steps)
    ok = np.array_equal(c_np[0, 0], want_local[0, 0]) if rank == 0 else True
    own_ms = own["s"] * 1e3
    # per-rank copy rates: the slowest rank's H2D / D2H stream-busy GB/s and the aggregate over ranks
    h2d_gbs, d2h_gbs = hs["h2d_bytes"] / (max(hs["h2d_ms"], 1e-6) * 1e-3) / 1e9, hs["d2h_bytes"] / (max(hs["d2h_ms"], 1e-6) * 1e-3) / 1e9
    e2e_host = {"h2d_GBs_min_rank": -cx.max_over_ranks(-h2d_gbs), "d2h_GBs_min_rank": -cx.max_over_ranks(-d2h_gbs),
                "h2d_GBs_sum": cx.sum_over_ranks(h2d_gbs), "d2h_GBs_sum": cx.sum_over_ranks(d2h_gbs),
                "kernel_ms_max_rank": cx.max_over_ranks(hs["kernel_ms"]), "groups": hs["groups"],
                "call_ms_fastest_rank": -cx.max_over_ranks(-own_ms), "call_ms_slowest_rank": cx.max_over_ranks(own_ms)}
    # pageable arrays (what a numpy caller of the reference holds): staged through the library's pinned ring
    pg_a, pg_b, pg_c = a_h.copy(), b_h.copy(), np.empty_like(a_h)
    pg_steps = max(3, min(K, 10))
    pg_s, pg_hs = e2e_leg(pg_a, pg_b, pg_c, pg_steps)
    ok = ok and (np.array_equal(pg_c[0, 0], want_local[0, 0]) if rank == 0 else True)
    assert pg_hs["staged"] == 1.0 and hs["staged"] == 0.0
    # ... and the same arrays handed to the driver as they are (cudaMemcpyAsync from pageable memory: the driver stages them
    # itself, one thread, and the three streams serialise) -- what the call did before the staging ring existed
    os.environ["GFHE_HOST_STAGING"] = "0"
    try:
        raw_s, _ = e2e_leg(pg_a, pg_b, pg_c, max(2, pg_steps // 2))
    finally:
        del os.environ["GFHE_HOST_STAGING"]
    # host memory copy rate with every rank copying at once (what staging and the DMA engines share)
    cx.barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        np.copyto(pg_c, pg_a)
    copy_gbs = 3 * 2 * pg_a.nbytes / (time.perf_counter() - t0) / 1e9
    host_copy = {"numpy_copy_GBs_per_rank_min": -cx.max_over_ranks(-copy_gbs), "numpy_copy_GBs_sum": cx.sum_over_ranks(copy_gbs),
                 "note": "np.copyto of the shard (read + write bytes), one thread per rank, all ranks at once"}
    assert cx.all_true(ok), "e2e (host-buffer) result is not bit-exact"
    del pg_a, pg_b, pg_c

    clocks.stop = True
    if prev_affinity:
        os.sched_setaffinity(0, prev_affinity)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    value = UNITS_TOTAL * K / (ms_total * 1e-3)
    dom = "k_row_polymul"
    dom_ms = kernels.get(dom, {}).get("ms_per_launch")
    step_kernel_ms = sum(k["ms_per_launch"] * k["launches_per_step"] for k in kernels.values()) or None
    # algorithmic bytes of ONE launch on one GPU: its shard, or half of it when the job is small enough for the two-stream
    # split (plan.cu split_run: two launches per pass, overlapping on two streams -- each one's duration then includes
    # sharing the SMs with the other half's kernels)
    dom_launches = max(1.0, round(kernels.get(dom, {}).get("launches_per_step", 1.0)))
    alg_launch = ALG_BYTES_POLYMUL * units_local / dom_launches
    achieved = alg_launch / (dom_ms * 1e-3) / 1e9 if dom_ms else None
    whole_gbs = ALG_BYTES_POLYMUL * UNITS_TOTAL / (ms_total / K * 1e-3) / 1e9      # aggregate over all GPUs
    ntt_value = UNITS_TOTAL * K / (ms_ntt * 1e-3)
    ntt_gbs = ALG_BYTES_NTT * UNITS_TOTAL / (ms_ntt / K * 1e-3) / 1e9

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, per_rep = cpu_polymul_rate(all_moduli, threads, reps=3, warm=1)
        cpu = {"value": v, "unit": "limb-poly-mul/s", "cores": threads, "kind": "port",
               "sample": "3 x full config-4 batch [32,16,65536] (512 limb-poly-muls each, %.2f s per batch) after 1 untimed warm-up batch"
                         % per_rep,
               "note": "oracle/ntt_oracle.c (Shoup/Harvey flavour, pthreads over limb-polynomials); the reference "
                       "repo itself has no NTT/poly-mul to time.  Same routine and warm-up as --impl reference"}

    # the element-wise family the reference DOES implement (arithmetic.py:3-13): its statement `(a + b) % MOD` /
    # `(a * b) % MOD` on dtype=object numpy arrays (the only mode that is exact for 60-bit moduli), one host core, on a
    # 2^20-residue sample of the same operands; next to it the GPU kernels' rate for the same ops
    cpu_vec = None
    if cpu is not None and ntt_domain:
        n_s = 1 << 20
        xa = a_h.reshape(-1)[:n_s].astype(object)
        xb = b_h.reshape(-1)[:n_s].astype(object)
        q0 = int(moduli[0])
        t0 = time.perf_counter(); _ = (xa + xb) % q0; t_add = time.perf_counter() - t0
        t0 = time.perf_counter(); _ = (xa * xb) % q0; t_mul = time.perf_counter() - t0
        cpu_vec = {"unit": "residue/s", "cores": 1, "kind": "reference statement",
                   "sample": "2^20 residues of operand a, b (limb 0), numpy dtype=object, statement of arithmetic.py:5,13",
                   "vec_add": n_s / t_add, "vec_mul": n_s / t_mul,
                   "gpu_vec_add": units_local * N / (ntt_domain["vec_add"]["ms_per_launch"] * 1e-3),
                   "gpu_pointwise_mul": units_local * N / (ntt_domain["mul_acc_terms1"]["ms_per_launch"] * 1e-3)}

    line = {
        "metric": "limb_poly_mul_per_s", "value": value, "unit": "limb-poly-mul/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_total / K,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(world),
        "full_rns_poly_mul_per_s": value / L_TOTAL,
        "value_with_gather": UNITS_TOTAL / (min(gather["step_with_gather_ms"], gather.get("step_with_overlapped_gather_ms", 1e30),
                                                     gather["peer_store"].get("step_with_gather_ms", 1e30), gather["peer_store"].get("step_with_overlapped_gather_ms", 1e30)) * 1e-3) if gather else value,
        "parity": parity,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak if achieved else None, "traffic": ncu_traffic(dom) if world == 1 else None,
                     "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_launch, "units_per_launch": units_local / dom_launches, "launches_per_step": dom_launches,
                     "ms_per_launch": dom_ms,
                     "share_of_step": dom_ms / step_kernel_ms if dom_ms and step_kernel_ms else None,
                     "whole_step": {"achieved": whole_gbs, "frac_of_aggregate": whole_gbs / (peak * world),
                                    "frac_of_8TBs_nominal_aggregate": whole_gbs / (8000.0 * world)},
                     "note": "60-bit limbs are bound by the 32-bit integer multiply pipe, not HBM (DESIGN.md section 3): "
                             "the ceiling of this formulation is ~16 % whole-step"},
        "kernels": kernels,
        "ntt": {"metric": "limb_ntt_per_s", "value": ntt_value, "unit": "limb-NTT/s", "ms_per_step": ms_ntt / K,
                "achieved_GBs": ntt_gbs, "frac_of_aggregate": ntt_gbs / (peak * world),
                "inverse": {"value": UNITS_TOTAL * K / (ms_intt * 1e-3), "ms_per_step": ms_intt / K,
                            "frac_of_aggregate": ALG_BYTES_NTT * UNITS_TOTAL / (ms_intt / K * 1e-3) / 1e9 / (peak * world)}},
        "e2e": {"value": UNITS_TOTAL / e2e_s, "unit": "limb-poly-mul/s", "steps": e2e_steps,
                "ms_per_step": e2e_s * 1e3,
                "h2d_bytes_per_step": int(2 * a_h.nbytes) * world, "d2h_bytes_per_step": int(a_h.nbytes) * world,
                "api": "gfhe_poly_mul_host, one call per rank on its shard (pinned host arrays; limb groups pipelined H2D/kernels/D2H)",
                "copies": e2e_host},
        "e2e_pageable": {"value": UNITS_TOTAL / pg_s, "unit": "limb-poly-mul/s", "steps": pg_steps, "ms_per_step": pg_s * 1e3,
                         "stage_in_ms": pg_hs["stage_in_ms"], "stage_out_ms": pg_hs["stage_out_ms"],
                         "unstaged_ms_per_step": raw_s * 1e3,
                         "api": "same call with plain (pageable) numpy arrays: staged through the library's pinned ring by host threads "
                                "(unstaged = GFHE_HOST_STAGING=0, the driver's own pageable path)"},
        "host_copy": host_copy,
        "gather": gather,
        "extra": {"weak": weak, "configs": configs, "config5_sweep": sweep,
                  "config5_note": "N=2^17 forward NTT, 60-bit limbs, L,B in {1..64}; each point uses min(n_gpus, L*B) GPUs (power of two, "
                                  "no transform is split across devices -- SURVEY 7 H8); frac = algorithmic bytes / time / (gpus x measured peak)"},
        "ntt_domain": ntt_domain,
        "moduli_30bit": (narrow if narrow_err is None else {"error": narrow_err}) if world == 1 else None,
        "cpu_baseline_vec_ops": cpu_vec,
        "gpu_launches": int(round(sum(k["launches_per_step"] for k in kernels.values()) * K)) if kernels else 3 * K,
        "clocks": clocks.summary(),
        "devices": {"rank0_device": cx.local, "map": "probe" if cx.device_probe else "identity (device = local rank)",
                    "probe": cx.device_probe},
        "cpu_baseline": cpu,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sweep", action="store_true", help="skip the config-5 N=2^17 grid (49 points)")
    args = ap.parse_args()
    claim_stdout()
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
