#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 drop-in for libsnark's Groth16 proving hot path.

Contract (one JSON line on stdout, rank 0):
    python bench.py --gpus N --steps K --warmup W            our arm
    python bench.py --impl reference --gpus N ...            the reference's CPU path (oracle/_ref)
    torchrun ... bench.py --gpus N ...                       N > 1: one rank per GPU

Workload = BASELINE.json configs[1]: alt_bn128 (BN254) G1 multi_exp over 2^20 distinct affine bases x
uniformly random Fr scalars (metric "G1 MSM Mpts/s"). One step = one whole MSM (scalar -> digits ->
bucket sort -> bucket accumulation -> bucket reduction -> window fold -> normalised point).
  value : bases AND scalars already resident in HBM when the timed region starts.
  e2e   : the reference-facing call (C-ABI b200_msm_resident = libff::multi_exp with the proving-key
          query cached on the device): scalars start in pinned HOST memory, the result ends on the host.
N > 1  : weak scaling by base range: every rank owns a 2^20-point shard of one
         N*2^20-point MSM, computes its per-window sums, the ranks all-gather them over NCCL and rank 0
         folds. value = N*2^20 points / max-over-ranks time.
Secondary metrics of BASELINE.json ride along: "ntt" (NTT elem/s @2^24, N = 1 only) and "groth16_prove" (Groth16 prove
ms @2^20 BN254: b200_groth16_prove at N = 1; at N > 1 the same fixed problem sharded over the N ranks by
libsnark_b200.DistributedProver, tools/prove_multi.py — strong scaling).
Inputs are synthetic (device-generated points, numpy scalars): nothing under oracle/ is touched by the
timed path; oracle/_ref is used for the cpu_baseline leg and for --impl reference only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CURVE = "alt_bn128"
LOG_N = 20
METRIC = "G1 MSM Mpts/s"
WORKLOAD = "alt_bn128 G1 multi_exp 2^20 distinct affine bases x random Fr scalars (BASELINE.json configs[1])"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--log-n", type=int, default=LOG_N)
    ap.add_argument("--no-ntt", action="store_true", help="skip the secondary NTT 2^24 measurement")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ntt-log-m", type=int, default=24)
    ap.add_argument("--no-prove", action="store_true", help="skip the Groth16 prove measurement")
    ap.add_argument("--prove-log-m", type=int, default=20)
    ap.add_argument("--no-tables", action="store_true",
                    help="do not precompute the [2^(c w)] P rows of the resident bases (b200_bases_precompute)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        return {"hbm_gbs": 6650.0, "fallback": True}


def msm_algorithmic_imads(n, c, W, n32=8, g2=False):
    """modmuls = 11 W N + 16 W 2^c (G1); one modmul = 2 n32^2 + n32 multiply-adds."""
    mm = 2 * n32 * n32 + n32
    modmuls = 11 * W * n + 16 * W * (1 << c)
    if g2:
        modmuls = 29 * W * n + 43 * W * (1 << c)
    return modmuls * mm


# ------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the unmodified reference's libff::multi_exp on the host cores
# ------------------------------------------------------------------------------------------------

def reference_msm_rate(log_n_full, steps, warmup, budget_s=60.0):
    """Times libff::multi_exp<G1, Fr, BDLO12_signed, special>(chunks = threads) from oracle/_ref on a
    bounded sample of the workload. Returns (Mpts/s, threads, sample description, ms per step)."""
    from oracle import refo
    ref = refo.RefOracle()
    # torchrun exports OMP_NUM_THREADS=1; the reference arm is entitled to every host core it can use
    ref.set_num_threads(len(os.sched_getaffinity(0)))
    threads = ref.num_threads()
    # size the sample from a small probe so that (steps + warmup) runs fit the budget
    probe_n = 1 << 14
    bases = ref.make_bases(CURVE, refo.G1, probe_n, 1)
    sc = ref.field_random(CURVE, refo.FR, 0, probe_n)
    t0 = time.time()
    ref.msm(CURVE, refo.G1, bases, sc, chunks=threads)
    per_pt = (time.time() - t0) / probe_n
    log_n = log_n_full
    while log_n > 14 and per_pt * (1 << log_n) * (steps + warmup) > budget_s:
        log_n -= 1
    n = 1 << log_n
    bases = ref.make_bases(CURVE, refo.G1, n, 1)
    sc = ref.field_random(CURVE, refo.FR, 0, n)
    for _ in range(warmup):
        ref.msm(CURVE, refo.G1, bases, sc, chunks=threads)
    t0 = time.time()
    for _ in range(steps):
        ref.msm(CURVE, refo.G1, bases, sc, chunks=threads)
    dt = (time.time() - t0) / steps
    sample = (f"libff::multi_exp<alt_bn128_G1, Fr, BDLO12_signed, special>(chunks={threads}) on 2^{log_n} of the "
              f"2^{log_n_full} points, {steps} timed runs after {warmup} warm-up, OMP threads={threads}")
    return n / dt / 1e6, threads, sample, dt * 1e3


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    rate, threads, sample, ms = reference_msm_rate(args.log_n, steps, warmup, budget_s=150.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "Mpts/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 limbs (254-bit modular)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "reference CPU path (oracle/_ref = unmodified libff), host cores only"},
        "cpu_baseline": {"value": rate, "unit": "Mpts/s", "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": rate, "unit": "Mpts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------

def run_b200_arm(args):
    import torch
    import torch.distributed as dist

    import libsnark_b200 as b

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    b.init(local_rank)
    dev = torch.device("cuda", local_rank)

    n = 1 << args.log_n
    K, Wm = args.steps, max(args.warmup, 3)
    # ---- synthetic inputs: this rank's shard of the global (world * n)-point MSM ----------------
    bases = b.make_bases(CURVE, b.G1, n, k0=1 + rank * n)
    rb = b.ResidentBases(CURVE, b.G1, bases)
    rb_plain = None if args.no_tables or world > 1 else b.ResidentBases(CURVE, b.G1, bases)   # the same key without tables
    del bases
    table_rows = 0 if args.no_tables else rb.precompute()
    scalars = b.random_scalars(CURVE, n, seed=1234 + rank)
    pinned = torch.empty((n, 4), dtype=torch.int64, pin_memory=True)
    pinned_np = pinned.numpy().view(np.uint64)
    pinned_np[:] = scalars
    dscal = b.DeviceBuffer(scalars.nbytes)
    dscal.upload(scalars)
    W = b.lib().b200_msm_num_windows(b.ALT_BN128, n)
    wbytes = b.lib().b200_msm_window_bytes(b.ALT_BN128, b.G1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    win_t = torch.empty(W * wbytes // 8, dtype=torch.int64, device=dev)
    gather_t = torch.empty(world * W * wbytes // 8, dtype=torch.int64, device=dev) if world > 1 else None

    def finish_multi():
        dist.all_gather_into_tensor(gather_t, win_t)
        if rank == 0:
            allw = gather_t.cpu().numpy().view(np.uint64)
            return b.fold_windows(CURVE, b.G1, n, allw, world)
        torch.cuda.synchronize()
        return None

    def step_resident():
        if world == 1:
            return rb.multi_exp(None, device_ptr=dscal.ptr)
        rb.window_sums_dev(win_t.data_ptr(), device_ptr=dscal.ptr)
        return finish_multi()

    def step_e2e():
        if world == 1:
            return rb.multi_exp(pinned_np)
        rb.window_sums_dev(win_t.data_ptr(), scalars=pinned_np)
        return finish_multi()

    def sync_all():
        torch.cuda.synchronize()
        b.lib().b200_sync()
        if world > 1:
            dist.barrier()

    def timed(fn, steps):
        """Per step: L2 flush (outside the timed span), barrier + sync, CUDA-event span on the library
        stream and host wall clock around the step, max over ranks."""
        ev_ms, wall_ms, stage = [], [], []
        for _ in range(steps):
            flush.zero_()
            sync_all()
            t0 = time.perf_counter()
            b.timer_start()
            fn()
            e = b.timer_stop()
            torch.cuda.synchronize()
            w = (time.perf_counter() - t0) * 1e3
            if world > 1:
                t = torch.tensor([e, w], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                e, w = t.tolist()
            ev_ms.append(e)
            wall_ms.append(w)
            stage.append(b.last_msm_timing())
        return ev_ms, wall_ms, stage

    for _ in range(Wm):
        r0 = step_resident()
        step_e2e()
    sampler = ClockSampler(local_rank)
    launches0 = b.launch_count()
    sampler.start()
    ev_ms, wall_ms, stage = timed(step_resident, K)
    clocks = sampler.stop()
    launches = b.launch_count() - launches0
    _, e2e_wall, _ = timed(step_e2e, K)
    r1 = step_e2e()
    if rank == 0 and r0 is not None and not np.array_equal(r0, r1):
        raise SystemExit("bench.py: resident and e2e paths disagree")

    plain = None
    if rb_plain is not None:
        for _ in range(Wm):
            rb_plain.multi_exp(None, device_ptr=dscal.ptr)
        _, pw, pst = timed(lambda: rb_plain.multi_exp(None, device_ptr=dscal.ptr), K)
        plain = {"value": n / (float(np.mean(pw)) * 1e-3) / 1e6, "unit": "Mpts/s", "ms_per_step": float(np.mean(pw)),
                 "stage_ms": {k: float(np.mean([s_[k] for s_ in pst])) for k in pst[0]},
                 "note": "the same MSM over the same resident key without b200_bases_precompute (17 windows, 17 bucket sets)"}
        rb_plain.free()

    # the host fold is part of the step, so the step time is the wall clock of the synchronous call
    ms_step = float(np.mean(wall_ms))
    total_pts = world * n
    value = total_pts / (ms_step * 1e-3) / 1e6
    e2e_ms = float(np.mean(e2e_wall))
    acc_ms = float(np.mean([s["accumulate"] for s in stage]))

    line = {
        "metric": METRIC, "value": value, "unit": "Mpts/s", "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32 limbs (254-bit modular)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "points_per_gpu": n, "parallelism": f"base-range shards x{world}",
                   "window_bits": int(b.lib().b200_msm_window_bits(b.ALT_BN128, n)), "windows": W,
                   "precomputed_table_rows": int(table_rows),
                   "tables": ("b200_bases_precompute: the resident key holds [2^(c w)] P for every window w (built once at "
                              "key upload, %d x 64 MiB), all windows share one bucket set; window_bits/windows above are "
                              "the table-free geometry the window-sum interface keeps" % table_rows) if table_rows else "off",
                   "l2": "256 MiB buffer written between timed iterations (outside the timed span)"},
        "ms_per_step_device_events": float(np.mean(ev_ms)),
        "stage_ms": {k: float(np.mean([s[k] for s in stage])) for k in stage[0]},
        "e2e": {"value": total_pts / (e2e_ms * 1e-3) / 1e6, "unit": "Mpts/s", "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": int(scalars.nbytes), "d2h_bytes_per_step": int(W * wbytes),
                "note": "b200_msm_resident: proving-key query cached on device, scalars from pinned host memory, "
                        "window sums back to the host, fold + normalise on the host"},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if plain is not None:
        line["without_tables"] = plain
    prove_multi_obj = None
    if world > 1 and not args.no_prove:
        # the whole prover sharded over the ranks (strong scaling of the fixed 2^20 problem); all ranks take part
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import prove_multi
        prove_multi_obj = prove_multi.run(b, torch, dist, CURVE, args.prove_log_m, max(3, min(args.steps, 10)),
                                          max(3, args.warmup), tables=not args.no_tables)
    if rank == 0:
        wide, lo, _ = b.imad_peak()
        c = line["config"]["window_bits"]
        # bucket accumulation only: (windows actually accumulated) * N mixed adds x 11 modmuls x MM(8)
        W_acc = table_rows if table_rows else W
        imads = 11 * W_acc * n * 136
        ach = imads / (acc_ms * 1e-3)
        line["roofline"] = {
            "bound": "int-pipe", "kernel": "k_accumulate", "achieved": ach / 1e12, "peak": wide / 1e12,
            "unit": "Tmad/s (32x32+64 multiply-adds)", "frac": ach / wide,
            # dram__bytes_read.sum + dram__bytes_write.sum of one k_accumulate launch at this workload,
            # profiles/r1_ncu_full_msm.md (ncu --set full); algorithmic bytes are 68 B per mixed add
            # dram__bytes_read.sum + dram__bytes_write.sum of one k_accumulate launch (ncu --set full): over the tables
            # profiles/r1s5_ncu_tables_msm.md, without them profiles/r1_ncu_full_msm.md
            "traffic": 2074.1e6 if table_rows else 525.8e6, "algorithmic_bytes": 68.0 * W_acc * n,
            "peak_source": "b200_imad_peak, measured in this run: chains of IMAD.WIDE.U32.X (the 32x32+64 multiply-add a "
                           "Montgomery product is made of; it holds the FMA-heavy pipe for two issue slots, 32 per clock "
                           "per SM). MEASURED_PEAKS.json has no integer figure.",
            "imad32_issue_peak": lo / 1e12, "frac_of_imad32_issue_peak": ach / lo,
            # sm__pipe_fmaheavy_cycles_active of the same captures
            "pipe_utilisation_ncu": 0.868 if table_rows else 0.839,
            "whole_msm_achieved": msm_algorithmic_imads(n, c, W) / (ms_step * 1e-3) / 1e12,
            "kernel_ms": acc_ms,
        }
        if world == 1 and not args.no_ntt:
            line["ntt"] = bench_ntt(b, torch, args.ntt_log_m, wide)
        if world == 1 and not args.no_prove:
            line["groth16_prove"] = bench_groth16(b, args.prove_log_m, max(3, min(args.steps, 10)), max(3, args.warmup),
                                                  tables=not args.no_tables)
        if prove_multi_obj is not None:
            line["groth16_prove"] = prove_multi_obj
        if world == 1 and not args.no_cpu_baseline:
            try:
                rate, threads, sample, _ = reference_msm_rate(args.log_n, 2, 1, budget_s=20.0)
                line["cpu_baseline"] = {"value": rate, "unit": "Mpts/s", "cores": threads, "kind": "reference",
                                        "sample": sample}
            except Exception as e:  # the checker library is optional for the timed arm
                line["cpu_baseline"] = {"value": None, "unit": "Mpts/s", "cores": 0, "kind": "reference",
                                        "sample": f"unavailable: {e}"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def bench_groth16(b, log_m, steps, warmup, tables=True):
    """BASELINE.json metric 'Groth16 prove ms @2^20 BN254': b200_groth16_prove (witness on the host in, proof on the host
    out) on a synthetic satisfiable R1CS of the shape of generate_r1cs_example_with_field_input with num_inputs = 10 and
    num_constraints = 2^log_m - 11, proving key = distinct device-generated points of the right
    sizes, resident on the device. Parity of this entry point with the reference prover (fixed r, s) is what
    tests/test_groth16_gpu.py::test_c_abi_prover_bit_exact checks; profiles/r1_groth16_prove.jsonl has the reference's time."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import synth_r1cs
    ni, m = 10, 1 << log_m
    t0 = time.time()
    sy = synth_r1cs.make(CURVE, m - ni - 1, ni)
    nv = sy["nv"]
    cs = b.R1CS(CURVE, sy["nc"], ni, nv, sy["row_ptr"], sy["col"], sy["val"])
    k = 1
    qs = []
    for grp, n in [(b.G1, nv + 1), (b.G2, nv + 1), (b.G1, nv + 1), (b.G1, m - 1), (b.G1, nv - ni)]:
        qs.append(b.make_bases(CURVE, grp, n, k0=k))
        k += n
    p1 = b.make_bases(CURVE, b.G1, 3, k0=k)
    p2 = b.make_bases(CURVE, b.G2, 2, k0=k + 3)
    pk = b.Groth16ProvingKey(CURVE, p1[0], p1[1], p2[0], p1[2], p2[1], qs[0], qs[1], qs[2], None, qs[3], qs[4], cs, m,
                             precompute=tables)
    setup_s = time.time() - t0
    rs = b.random_scalars(CURVE, 2, seed=99)
    # the witness lies in page-locked host memory, as the e2e contract says (the library then skips its staging copy)
    import torch
    aux_pinned = torch.empty(sy["auxiliary"].shape, dtype=torch.int64, pin_memory=True)
    aux_np = aux_pinned.numpy().view(np.uint64)
    aux_np[:] = sy["auxiliary"]
    sy = dict(sy, auxiliary=aux_np)
    proof = None
    for _ in range(warmup):
        proof, _ = pk.prove(sy["primary"], sy["auxiliary"], rs[0], rs[1])
    times, stages = [], []
    for _ in range(steps):
        t0 = time.perf_counter()
        p, st = pk.prove(sy["primary"], sy["auxiliary"], rs[0], rs[1])
        times.append((time.perf_counter() - t0) * 1e3)
        stages.append(st)
        assert np.array_equal(p, proof), "prover is not deterministic for fixed r, s"
    ms = float(np.mean(times))
    keys = stages[0].keys()
    out = {"metric": f"Groth16 prove ms @2^{log_m} BN254", "value": ms, "unit": "ms", "higher_is_better": False,
           "steps": steps, "warmup": warmup, "min_ms": float(np.min(times)),
           "stage_ms": {kk: float(np.mean([s_[kk] for s_ in stages])) for kk in keys},
           "sizes": {"num_constraints": sy["nc"], "num_inputs": ni, "num_variables": nv, "domain": m,
                     "r1cs_nnz": int(sy["col"].shape[0])},
           "h2d_bytes_per_step": int((nv + 1) * 32), "d2h_bytes_per_step": int(5 * 17 * 128 + 17 * 256),
           "setup_s": setup_s, "precomputed_tables": bool(tables),
           "note": "wall clock around b200_groth16_prove: witness (host) -> H2D -> sparse R1CS evaluation -> 7 NTTs -> "
                   "A, B(G2+G1), H, L MSMs on device-resident scalars -> window sums D2H -> host fold and proof assembly; "
                   "proving key and constraint system resident on the device (uploaded once, outside the timed region)"}
    pk.free()
    cs.free()
    return out


def bench_ntt(b, torch, log_m, wide_peak):
    """Secondary metric 'NTT elem/s @2^24': basic_radix2_domain<alt_bn128_Fr>::FFT, data resident in HBM
    (value) and through the host-buffer C-ABI call b200_ntt (e2e)."""
    m = 1 << log_m
    data = b.random_scalars(CURVE, m, seed=77)
    buf = b.DeviceBuffer(data.nbytes)
    buf.upload(data)
    import ctypes as C
    L = b.lib()
    for _ in range(3):
        assert L.b200_ntt_dev(b.ALT_BN128, b.FFT, C.c_void_p(buf.ptr), m, None) == 0
    steps = 5
    b.timer_start()
    for _ in range(steps):
        assert L.b200_ntt_dev(b.ALT_BN128, b.FFT, C.c_void_p(buf.ptr), m, None) == 0
    ms = b.timer_stop() / steps
    buf.free()
    # e2e: the reference-facing call (b200_ntt = basic_radix2_domain<Fr>::FFT, in place on a HOST vector), here a
    # page-locked one: H2D of the 2^24 elements, the transform, D2H of the result, all inside the timed call
    pinned = torch.empty((m, 4), dtype=torch.int64, pin_memory=True)
    pn = pinned.numpy().view(np.uint64)
    pn[:] = data
    assert L.b200_ntt(b.ALT_BN128, b.FFT, C.c_void_p(pn.ctypes.data), m, None) == 0
    e2e = []
    for _ in range(3):
        t0 = time.perf_counter()
        assert L.b200_ntt(b.ALT_BN128, b.FFT, C.c_void_p(pn.ctypes.data), m, None) == 0
        e2e.append((time.perf_counter() - t0) * 1e3)
    e2e_ms = float(np.mean(e2e))
    del pinned, pn
    peaks = measured_peaks()
    gbs = 2 * m * 32 / (ms * 1e-3) / 1e9
    return {"metric": "NTT elem/s", "log_m": log_m, "value": m / (ms * 1e-3), "unit": "elem/s", "ms": ms,
            "e2e": {"value": m / (e2e_ms * 1e-3), "unit": "elem/s", "ms": e2e_ms,
                    "h2d_bytes_per_step": int(data.nbytes), "d2h_bytes_per_step": int(data.nbytes)},
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
                         "frac": gbs / peaks.get("hbm_gbs"),
                         # dram__bytes_read + dram__bytes_write of one k_ntt_pass launch without scale tables (0.537 + 0.490 GB,
                         # profiles/r1_ncu_ntt_pass.md) x the 3 passes of a 2^24 transform; scaled by size for other log_m
                         "traffic": 3 * (0.537e9 + 0.4896e9) * (m / float(1 << 24)) * (-(-log_m // 8) / 3.0),
                         "algorithmic_bytes": 2 * m * 32,
                         "int_pipe_note": "(m/2) log2(m) modmuls x 136 multiply-adds against the IMAD.WIDE peak; see DESIGN.md §5"},
            "modmul_Tmad_s": (m // 2) * log_m * 136 / (ms * 1e-3) / 1e12,
            "int_pipe_frac": (m // 2) * log_m * 136 / (ms * 1e-3) / wide_peak}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
