#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 drop-in for libsnark's Groth16 proving hot path.

Contract (one JSON line on stdout, rank 0):
    python bench.py --gpus N --steps K --warmup W            our arm
    python bench.py --impl reference --gpus N ...            the reference's own CPU prover (oracle/_ref)
    torchrun ... bench.py --gpus N ...                       N > 1: one rank per GPU
    python bench.py --config 4|5 ...                         BASELINE.json configs[3] / configs[4] at full size

Headline = BASELINE.json's first-named metric, "Groth16 prove ms @2^20 BN254": r1cs_gg_ppzksnark_prover
(r1cs_gg_ppzksnark.tcc:460-603) on alt_bn128 over the R1CS of generate_r1cs_example_with_field_input with
num_inputs = 10, num_constraints = 2^20 - 11 (domain 2^20), r and s fixed.
  value : b200_groth16_prove_resident — witness already in HBM, proving key resident, proof on the host.
  e2e   : b200_groth16_prove — witness in pinned HOST memory, H2D inside the timed call, proof on the host.
Both arms run on the SAME inputs: the example and a proving key of distinct affine points are built by the
unmodified reference (oracle/_ref, ref_harness.cpp::setup) and handed to the GPU prover; after the timed region the
reference prover runs once on them (that run is the cpu_baseline) and the three proof points are compared
("parity_checked"). Without oracle/_ref the bench still runs on a self-made synthetic key and says parity_checked false.
N > 1: the same fixed 2^20 problem sharded over the ranks (libsnark_b200.DistributedProver): STRONG scaling.
Secondary objects of the same line: "msm" (G1 MSM Mpts/s, BASELINE configs[1]; weak scaling at N > 1), "ntt" (NTT
elem/s @2^24, configs[2], N = 1), "config4" (bls12_377 prove at 2^22 incl. G2) and "sweep" (configs[4] points).
Nothing under oracle/ is touched inside a timed GPU region; oracle/_ref provides inputs, the parity check, the
cpu_baseline legs and the --impl reference arm.
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
sys.path.insert(0, os.path.join(ROOT, "tools"))

CURVE = "alt_bn128"
NI = 10
METRIC = "Groth16 prove ms @2^20 BN254"
WORKLOAD = ("r1cs_gg_ppzksnark (Groth16) prove on alt_bn128, 2^20-constraint-domain synthetic R1CS from "
            "generate_r1cs_example_with_field_input (num_inputs 10, num_constraints 2^20 - 11), fixed r/s "
            "(BASELINE.json metric; configs[0] at the size the metric names)")
MSM_WORKLOAD = "alt_bn128 G1 multi_exp 2^20 distinct affine bases x random Fr scalars (BASELINE.json configs[1])"
MM = {8: 136, 12: 300, 24: 1176}       # multiply-adds per Montgomery product, SURVEY.md §8(d)
N32_FQ = {"alt_bn128": 8, "bls12_377": 12, "bw6_761": 24}
N32_FR = {"alt_bn128": 8, "bls12_377": 8, "bw6_761": 12}
FR_BITS = {"alt_bn128": 254, "bls12_377": 253, "bw6_761": 377}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=1, choices=[1, 4, 5],
                    help="1: the headline line (default); 4: bls12_377 prove at 2^22 with the reference at full size; "
                         "5: the MSM size sweep on bls12_377 / bw6_761 G1")
    ap.add_argument("--log-n", type=int, default=20, help="MSM size (secondary metric)")
    ap.add_argument("--prove-log-m", type=int, default=20)
    ap.add_argument("--ntt-log-m", type=int, default=24)
    ap.add_argument("--config4-log-m", type=int, default=22)
    ap.add_argument("--sweep-logs", type=int, nargs="+", default=None)
    ap.add_argument("--no-ntt", action="store_true")
    ap.add_argument("--no-msm", action="store_true")
    ap.add_argument("--no-config4", action="store_true")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip every reference run (no parity flags either)")
    ap.add_argument("--no-tables", action="store_true",
                    help="do not precompute the [2^(c w)] P rows of the resident bases (b200_bases_precompute)")
    return ap.parse_args()


def host_cores():
    return len(os.sched_getaffinity(0))


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


def ncu_constants():
    """dram bytes / pipe utilisation of the dominant kernels from the committed ncu --set full captures
    (profiles/ncu_constants.json, written by hand from the capture summaries next to it). Not measured in this run:
    the keys that carry them say `_ncu`."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_constants.json")))
    except (OSError, ValueError):
        return {}


def msm_modmuls(n, c, W, reduce_sets, g2=False):
    """SURVEY.md §8(d) accounting with the geometry actually run: 11 modmuls per mixed add (W n of them), 16 per bucket
    of each bucket set that is reduced (W sets without tables, one over tables); x 29/11 and 43/16 for G2 over Fq2."""
    if g2:
        return 29 * W * n + 43 * reduce_sets * (1 << c)
    return 11 * W * n + 16 * reduce_sets * (1 << c)


def load_ref():
    """oracle/_ref (the unmodified reference compiled by oracle/Makefile) with every host core, or None."""
    try:
        from oracle import refo
        ref = refo.RefOracle()
        ref.set_num_threads(host_cores())   # torchrun exports OMP_NUM_THREADS=1; the reference may use every core
        return ref
    except (OSError, FileNotFoundError, ImportError, AssertionError):
        return None


# ------------------------------------------------------------------------------------------------
# inputs: one R1CS example + proving key, shared by both arms
# ------------------------------------------------------------------------------------------------

PROVE_FIELDS = ["row_ptr", "col", "val", "points", "A", "Bg2", "Bg1", "Bidx", "H", "L", "primary", "auxiliary", "rs"]


def prove_inputs_from_reference(ref, curve, log_m):
    """The example and a synthetic proving key (queries of the right sizes filled with distinct affine points,
    oracle/ref_harness.cpp::setup with real_keys = 0) built by the reference itself; returns (arrays, fixture).
    The prover's work and the CPU/GPU parity do not depend on the key being a real Groth16 key."""
    from oracle import refo
    m = 1 << log_m
    fx = ref.groth16_setup(curve, m - NI - 1, NI, real_keys=False)
    sz, g2l = fx.sizes, fx.i["g2_limbs"]
    row_ptr, col, val = fx.r1cs_csr()
    pts = fx.pk_points()
    d = {"nc": sz["num_constraints"], "ni": sz["num_inputs"], "nv": sz["num_variables"], "m": sz["m"],
         "row_ptr": row_ptr, "col": col, "val": val,
         "points": np.concatenate([np.asarray(p, dtype=np.uint64).reshape(-1) for p in pts]),
         "A": fx.A_query, "Bg2": np.ascontiguousarray(fx.B_values[:, :g2l]),
         "Bg1": np.ascontiguousarray(fx.B_values[:, g2l:]), "Bidx": fx.B_indices.astype(np.uint32),
         "H": fx.H_query, "L": fx.L_query, "primary": fx.primary_input, "auxiliary": fx.auxiliary_input,
         "rs": ref.field_random(curve, refo.FR, 0xABCDEF, 2), "source": "reference"}
    return d, fx


def prove_inputs_synthetic(b, curve, log_m):
    """Fallback when oracle/_ref is absent: tools/synth_r1cs.py + device-generated multiples of the generator."""
    import synth_r1cs
    m = 1 << log_m
    sy = synth_r1cs.make(curve, m - NI - 1, NI)
    nv = sy["nv"]
    k, qs = 1, []
    for grp, n in [(b.G1, nv + 1), (b.G2, nv + 1), (b.G1, nv + 1), (b.G1, m - 1), (b.G1, nv - NI)]:
        qs.append(b.make_bases(curve, grp, n, k0=k))
        k += n
    p1 = b.make_bases(curve, b.G1, 3, k0=k)
    p2 = b.make_bases(curve, b.G2, 2, k0=k + 3)
    pts = [p1[0], p1[1], p2[0], p1[2], p2[1]]
    return {"nc": sy["nc"], "ni": NI, "nv": nv, "m": m, "row_ptr": sy["row_ptr"], "col": sy["col"], "val": sy["val"],
            "points": np.concatenate([np.asarray(p, dtype=np.uint64).reshape(-1) for p in pts]),
            "A": qs[0], "Bg2": qs[1], "Bg1": qs[2], "Bidx": None, "H": qs[3], "L": qs[4],
            "primary": sy["primary"], "auxiliary": sy["auxiliary"], "rs": b.random_scalars(curve, 2, seed=99),
            "source": "synthetic (no oracle/_ref on this box)"}


def split_points(b, curve, flat):
    g1, g2 = b.sizeof_group(curve, b.G1) // 8, b.sizeof_group(curve, b.G2) // 8
    cuts = np.cumsum([0, g1, g1, g2, g1, g2])
    return [np.ascontiguousarray(flat[cuts[i]:cuts[i + 1]]) for i in range(5)]


def broadcast_inputs(d, torch, dist, rank, dev):
    """Rank 0 built the inputs; the other ranks receive them over NCCL (setup, outside every timed region)."""
    meta = [None]
    if rank == 0:
        meta[0] = {k: (None if d.get(k) is None else (d[k].shape, str(d[k].dtype))) for k in PROVE_FIELDS}
        meta[0]["_scalars"] = {k: d[k] for k in ("nc", "ni", "nv", "m", "source")}
    dist.broadcast_object_list(meta, src=0)
    out = dict(meta[0]["_scalars"]) if rank != 0 else d
    for k in PROVE_FIELDS:
        spec = meta[0][k]
        if spec is None:
            out[k] = None
            continue
        shape, dt = spec
        nbytes = int(np.prod(shape)) * np.dtype(dt).itemsize
        if rank == 0:
            t = torch.from_numpy(np.ascontiguousarray(d[k]).view(np.uint8).reshape(-1)).to(dev)
        else:
            t = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        dist.broadcast(t, src=0)
        if rank != 0:
            out[k] = t.cpu().numpy().view(np.dtype(dt)).reshape(shape)
        del t
    return out


# ------------------------------------------------------------------------------------------------
# reference legs (CPU): the unmodified reference from oracle/_ref on the host cores
# ------------------------------------------------------------------------------------------------

def reference_prove(ref, fx, rs, steps, warmup, budget_s):
    """r1cs_gg_ppzksnark_prover (r1cs_gg_ppzksnark.tcc:464-603, fixed r/s harness) on the fixture.
    Returns (proof, ms list, stage dict, steps done, warm-ups done)."""
    proof, times, stages = None, [], None
    t_start = time.time()
    done_w = 0
    for _ in range(warmup):
        proof, stages = fx.prove(rs[0], rs[1])
        done_w += 1
        if time.time() - t_start > budget_s * 0.4:
            break
    for k in range(steps):
        t0 = time.perf_counter()
        proof, stages = fx.prove(rs[0], rs[1])
        times.append((time.perf_counter() - t0) * 1e3)
        per = (time.time() - t_start) / (done_w + k + 1)
        if time.time() - t_start + per > budget_s:
            break
    return proof, times, stages, len(times), done_w


def reference_msm_rate(ref, curve, bases, scalars, steps, warmup):
    from oracle import refo
    threads = ref.num_threads()
    out = None
    for _ in range(warmup):
        out = ref.msm(curve, refo.G1, bases, scalars, chunks=threads)
    t0 = time.perf_counter()
    for _ in range(steps):
        out = ref.msm(curve, refo.G1, bases, scalars, chunks=threads)
    dt = (time.perf_counter() - t0) / steps
    return out, scalars.shape[0] / dt / 1e6, dt * 1e3


def reference_ntt(ref, curve, data):
    """basic_radix2_domain<Fr>::FFT (basic_radix2_domain.tcc:54-60), one run: (output, ms)."""
    from oracle import refo
    t0 = time.perf_counter()
    out = ref.fft(curve, refo.FFT, data)
    return out, (time.perf_counter() - t0) * 1e3


def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    ref = load_ref()
    if ref is None:
        emit({"impl": "reference", "unavailable": "oracle/_ref/libref_oracle.so is not built"})
        return
    from oracle import refo
    threads = ref.num_threads()
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    if args.config == 5:
        return run_sweep(args, None, ref, reference_only=True)
    curve, log_m = (CURVE, args.prove_log_m) if args.config == 1 else ("bls12_377", args.config4_log_m)
    t0 = time.time()
    m = 1 << log_m
    fx = ref.groth16_setup(curve, m - NI - 1, NI, real_keys=False)
    rs = ref.field_random(curve, refo.FR, 0xABCDEF, 2)
    setup_s = time.time() - t0
    _, times, stages, done, done_w = reference_prove(ref, fx, rs, steps, min(warmup, 1), budget_s=200.0)
    ms = float(np.mean(times))
    sample = (f"r1cs_gg_ppzksnark_prover<{curve}> (r1cs_gg_ppzksnark.tcc:464-603, r/s fixed) on the whole 2^{log_m} "
              f"problem, {done} timed runs after {done_w} warm-up (of the {steps}/{warmup} requested: bounded to ~200 s), "
              f"OMP threads={threads}")
    line = {
        "impl": "reference", "metric": METRIC if args.config == 1 else f"Groth16 prove ms @2^{log_m} {curve}",
        "value": ms, "unit": "ms", "n_gpus": args.gpus, "steps": done, "warmup": done_w,
        "steps_requested": steps, "ms_per_step": ms, "higher_is_better": False, "scaling": "strong",
        "vs_baseline": None, "dtype": "u32 limbs (254-bit modular)", "data": "synthetic",
        "config": {"workload": WORKLOAD if args.config == 1 else config4_workload(log_m),
                   "note": "reference CPU path (oracle/_ref = unmodified libsnark/libff/libfqfft), host cores only"},
        "stage_ms": stages, "setup_s": setup_s,
        "cpu_baseline": {"value": ms, "unit": "ms", "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": ms, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    fx.close()
    if args.config == 1 and not args.no_msm:
        n = 1 << args.log_n
        bases = ref.make_bases(CURVE, refo.G1, n, 1)
        sc = bench_scalars(CURVE, n, 1234)
        _, rate, mms = reference_msm_rate(ref, CURVE, bases, sc, 3, 1)
        line["msm"] = {"metric": "G1 MSM Mpts/s", "value": rate, "unit": "Mpts/s", "ms": mms, "cores": threads,
                       "sample": f"libff::multi_exp<alt_bn128_G1, Fr, BDLO12_signed, special>(chunks={threads}) on all 2^{args.log_n} "
                                 f"points, 3 timed runs after 1 warm-up"}
        del bases, sc
    if args.config == 1 and not args.no_ntt:
        mm_ = 1 << args.ntt_log_m
        data = bench_scalars(CURVE, mm_, 77)
        _, nms = reference_ntt(ref, CURVE, data)
        line["ntt"] = {"metric": "NTT elem/s", "log_m": args.ntt_log_m, "value": mm_ / (nms * 1e-3), "unit": "elem/s",
                       "ms": nms, "cores": threads,
                       "sample": f"basic_radix2_domain<alt_bn128_Fr>::FFT (basic_radix2_domain.tcc:54-60) at 2^{args.ntt_log_m}, "
                                 f"one run, no warm-up (SURVEY.md §8d: run once and say so)"}
    emit(line)


def config4_workload(log_m):
    return (f"Groth16 prove on bls12_377 at 2^{log_m} constraints incl. G2 MSM (BASELINE.json configs[3]; bls12_381 is not "
            f"in the reference tree, SURVEY.md F3), MSMs sharded by base range over the GPUs")


_R = {"alt_bn128": 0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001,
      "bls12_377": 0x12ab655e9a2ca55660b44d1e5c37b00159aa76fed00000010a11800000000001,
      "bw6_761": 0x1ae3a4617c510eac63b05c06ca1493b1a22d9f300f5138f1ef3622fba094800170b5d44300000008508c00000000001}


def bench_scalars(curve, n, seed):
    """n uniformly random Fr residues, identical in both arms (numpy; mirrors libsnark_b200.random_scalars so the
    reference arm does not import the product)."""
    r = _R[curve]
    L = (r.bit_length() + 63) // 64
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 64, size=(n, L), dtype=np.uint64)
    a[:, L - 1] %= np.uint64(r >> (64 * (L - 1)))
    return a


# ------------------------------------------------------------------------------------------------
# our arm: shared plumbing
# ------------------------------------------------------------------------------------------------

class Ctx:
    """torch / torch.distributed / library handles of this rank."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        import libsnark_b200 as b
        self.torch, self.dist, self.b, self.args = torch, dist, b, args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
        torch.cuda.set_device(self.local_rank)
        if self.world > 1:
            # NCCL on high-priority streams: its few CTAs must not queue behind the MSMs' machine-filling kernels
            opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank), pg_options=opts)
        b.init(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2
        self.ref = None if (args.no_cpu_baseline or self.rank != 0) else load_ref()
        self.tables = not args.no_tables

    def sync_all(self):
        self.torch.cuda.synchronize()
        self.b.lib().b200_sync()
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, vals):
        if self.world == 1:
            return list(vals)
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()

    def timed(self, fn, steps, after=None):
        """Per step: L2 flush (outside the timed span), barrier + sync on both sides, CUDA-event span on the library
        stream and host wall clock around the synchronous call, max over ranks. Returns (event ms, wall ms, extras)."""
        b = self.b
        ev_ms, wall_ms, extra = [], [], []
        for _ in range(steps):
            self.flush.zero_()
            self.sync_all()
            t0 = time.perf_counter()
            b.timer_start()
            r = fn()
            e = b.timer_stop()
            self.torch.cuda.synchronize()
            w = (time.perf_counter() - t0) * 1e3
            e, w = self.max_over_ranks([e, w])
            ev_ms.append(e)
            wall_ms.append(w)
            extra.append(after(r) if after else None)
        return ev_ms, wall_ms, extra

    def pinned_copy(self, a):
        a = np.ascontiguousarray(a)
        t = self.torch.empty(a.shape, dtype=self.torch.int64, pin_memory=True)
        v = t.numpy().view(np.uint64)
        v[:] = a
        return t, v


def mean_stages(stages):
    return {k: float(np.mean([s[k] for s in stages])) for k in stages[0]}


# ------------------------------------------------------------------------------------------------
# Groth16 prove (headline; also config 4 on bls12_377)
# ------------------------------------------------------------------------------------------------

def prove_algorithmic(ctx, curve, d, sets):
    """Algorithmic multiply-adds of one proof in SURVEY.md §8(d)'s accounting, with the window geometry the kernels
    actually ran: five MSMs + seven transforms of the witness map."""
    b, L = ctx.b, ctx.b.lib()
    cid = b._cid(curve)
    nv, ni, m = d["nv"], d["ni"], d["m"]
    sizes = [nv + 1, d["Bg2"].shape[0] if d.get("Bg2") is not None else nv + 1, None, m - 1, nv - ni]
    sizes[2] = sizes[1]
    groups = [False, True, False, False, False]
    mm_q, mm_r = MM[N32_FQ[curve]], MM[N32_FR[curve]]
    total, per = 0, []
    for n, g2, rb in zip(sizes, groups, sets):
        rows = L.b200_bases_table_rows(rb.handle) if rb is not None else 0
        if rows:
            c, W, rs = L.b200_bases_table_window_bits(rb.handle), rows, 1
        else:
            c, W = L.b200_msm_window_bits(cid, n), L.b200_msm_num_windows(cid, n)
            rs = W
        x = msm_modmuls(n, c, W, rs, g2) * mm_q
        per.append({"n": int(n), "c": int(c), "windows": int(W), "bucket_sets_reduced": int(rs), "imads": float(x)})
        total += x
    lg = (m - 1).bit_length()
    ntt = 7 * (m // 2) * lg * mm_r if m & (m - 1) == 0 else 0
    return float(total), float(ntt), per


def bench_prove_single(ctx, curve, log_m, steps, warmup, want_parity=True, ref_budget_s=60.0):
    """One-GPU prove: value (witness resident), e2e (witness on the host), stage split, parity with the reference
    prover on the same inputs and its time as cpu_baseline."""
    b, args = ctx.b, ctx.args
    t0 = time.time()
    fx = None
    if ctx.ref is not None:
        d, fx = prove_inputs_from_reference(ctx.ref, curve, log_m)
    else:
        d = prove_inputs_synthetic(b, curve, log_m)
    cs = b.R1CS(curve, d["nc"], d["ni"], d["nv"], d["row_ptr"], d["col"], d["val"])
    pts = split_points(b, curve, d["points"])
    pk = b.Groth16ProvingKey(curve, *pts, d["A"], d["Bg2"], d["Bg1"], d["Bidx"], d["H"], d["L"], cs, d["m"],
                             precompute=ctx.tables)
    setup_s = time.time() - t0
    rs = d["rs"]
    _, aux = ctx.pinned_copy(d["auxiliary"])   # page-locked, as the e2e contract says: the library skips its staging copy
    _keep = _
    proof = None
    for _i in range(warmup):
        proof, _st = pk.prove(d["primary"], aux, rs[0], rs[1])
    # e2e: host witness in, proof out
    _, e2e_wall, e2e_st = ctx.timed(lambda: pk.prove(d["primary"], aux, rs[0], rs[1]), steps, after=lambda r: r[1])
    # value: witness resident
    pk.upload_witness(d["primary"], aux)
    for _i in range(2):
        pk.prove_resident(rs[0], rs[1])
    sampler = ClockSampler(ctx.local_rank)
    launches0 = b.launch_count()
    sampler.start()
    ev, wall, st = ctx.timed(lambda: pk.prove_resident(rs[0], rs[1]), steps,
                             after=lambda r: (r[1], bool(np.array_equal(r[0], proof))))
    clocks = sampler.stop()
    launches = b.launch_count() - launches0
    assert all(s[1] for s in st), "prover is not deterministic for fixed r, s"
    ms, e2e_ms = float(np.mean(wall)), float(np.mean(e2e_wall))
    msm_imads, ntt_imads, per = prove_algorithmic(ctx, curve, d, pk.sets)
    fr = b.sizeof_fr(curve)
    out = {"value": ms, "unit": "ms", "min_ms": float(np.min(wall)), "ms_device_events_slot0": float(np.mean(ev)),
           "steps": steps, "warmup": warmup,
           "stage_ms": mean_stages([s[0] for s in st]),
           "e2e": {"value": e2e_ms, "unit": "ms", "min_ms": float(np.min(e2e_wall)),
                   "h2d_bytes_per_step": int((d["nv"] + 1) * fr),
                   "d2h_bytes_per_step": int(sum(b.lib().b200_msm_num_windows(b._cid(curve), max(1, r_.n)) *
                                                 b.lib().b200_msm_window_bytes(b._cid(curve), r_.group) for r_ in pk.sets)),
                   "stage_ms": mean_stages(e2e_st),
                   "note": "b200_groth16_prove: witness in page-locked host memory -> H2D -> sparse R1CS evaluation -> 7 NTTs -> "
                           "A, B(G2+G1), H, L MSMs -> window sums D2H -> host fold + proof assembly; key resident"},
           "sizes": {"num_constraints": d["nc"], "num_inputs": d["ni"], "num_variables": d["nv"], "domain": d["m"],
                     "r1cs_nnz": int(d["col"].shape[0])},
           "gpu_launches": int(launches), "clocks": clocks, "setup_s": setup_s, "precomputed_tables": bool(ctx.tables),
           "inputs": d["source"], "msm_geometry": per,
           "algorithmic_imads": {"msm": msm_imads, "ntt": ntt_imads}}
    parity, cpu = False, None
    if fx is not None and want_parity:
        want, times, ref_st, done, done_w = reference_prove(ctx.ref, fx, rs, 1, 0, ref_budget_s)
        parity = bool(np.array_equal(want, proof))
        if not parity:
            raise SystemExit("bench.py: GPU proof differs from the reference prover's on the same inputs")
        cpu = {"value": float(np.mean(times)), "unit": "ms", "cores": ctx.ref.num_threads(), "kind": "reference",
               "stage_ms": ref_st,
               "sample": f"r1cs_gg_ppzksnark_prover<{curve}> (r1cs_gg_ppzksnark.tcc:464-603, r/s fixed) on the same 2^{log_m} "
                         f"inputs, one run, no warm-up, OMP threads={ctx.ref.num_threads()}; its proof is the parity check"}
    out["parity_checked"] = parity
    out["parity"] = ("the three proof points equal the reference prover's, bit for bit, on the same example, key, r and s"
                     if parity else "not checked against the reference in this run (" + d["source"] + ")")
    if cpu:
        out["cpu_baseline"] = cpu
    if fx is not None:
        fx.close()
    pk.free()
    cs.free()
    return out


def bench_prove_sharded(ctx, curve, log_m, steps, warmup, use_reference_inputs):
    """N > 1: the fixed problem cut over the ranks by base range (DistributedProver). With use_reference_inputs rank 0
    builds example + key with the reference and broadcasts them, and the N-rank proof is compared with the reference
    prover's; otherwise every rank generates its own shard on the device and rank 0 compares with a one-GPU proof."""
    b, dist, torch = ctx.b, ctx.dist, ctx.torch
    world, rank = ctx.world, ctx.rank
    t0 = time.time()
    fx, d = None, None
    have_ref = [bool(use_reference_inputs and ctx.ref is not None)] if rank == 0 else [None]
    dist.broadcast_object_list(have_ref, src=0)
    m = 1 << log_m
    if have_ref[0]:
        if rank == 0:
            d, fx = prove_inputs_from_reference(ctx.ref, curve, log_m)
        d = broadcast_inputs(d, torch, dist, rank, ctx.dev)
        cs = b.R1CS(curve, d["nc"], d["ni"], d["nv"], d["row_ptr"], d["col"], d["val"])
        pts = split_points(b, curve, d["points"])
        pk = b.Groth16ProvingKeyShard(curve, *pts, d["A"], d["Bg2"], d["Bg1"], d["Bidx"], d["H"], d["L"], cs, d["m"], rank,
                                      world, sliced=False, precompute=ctx.tables)
        totals = k0s = None
    else:
        import synth_r1cs
        sy = synth_r1cs.make(curve, m - NI - 1, NI)
        nv = sy["nv"]
        d = {"nc": sy["nc"], "ni": NI, "nv": nv, "m": m, "primary": sy["primary"], "auxiliary": sy["auxiliary"],
             "rs": b.random_scalars(curve, 2, seed=99), "source": "synthetic: every rank generates its own shard on the device"}
        cs = b.R1CS(curve, sy["nc"], NI, nv, sy["row_ptr"], sy["col"], sy["val"])
        totals = [(b.G1, nv + 1), (b.G2, nv + 1), (b.G1, nv + 1), (b.G1, m - 1), (b.G1, nv - NI)]
        k0s, k = [], 1
        for _, n in totals:
            k0s.append(k)
            k += n
        shards = []
        for (grp, n), k0 in zip(totals, k0s):
            f, c = b.shard_range(n, world, rank)
            shards.append(b.make_bases(curve, grp, c, k0=k0 + f))
        p1 = b.make_bases(curve, b.G1, 3, k0=k)
        p2 = b.make_bases(curve, b.G2, 2, k0=k + 3)
        pts = (p1[0], p1[1], p2[0], p1[2], p2[1])
        pk = b.Groth16ProvingKeyShard(curve, *pts, shards[0], shards[1], shards[2], None, shards[3], shards[4], cs, m, rank,
                                      world, sliced=True, B_total=nv + 1, precompute=ctx.tables)
        del shards
    prover = b.DistributedProver(pk)
    setup_s = time.time() - t0
    rs = d["rs"]
    aux = d["auxiliary"]
    if rank == 0:
        _keep, aux = ctx.pinned_copy(aux)
    proof = None
    for _ in range(warmup):
        proof = prover.prove(d["primary"], aux, rs[0], rs[1])
    sampler = ClockSampler(ctx.local_rank)
    launches0 = b.launch_count()
    sampler.start()
    _, wall, st = ctx.timed(lambda: prover.prove(d["primary"], aux, rs[0], rs[1]), steps,
                            after=lambda p: (dict(prover.last_stages), rank != 0 or bool(np.array_equal(p, proof))))
    clocks = sampler.stop()
    launches = b.launch_count() - launches0
    assert all(s[1] for s in st), "prover is not deterministic for fixed r, s"
    ms = float(np.mean(wall))
    fr = b.sizeof_fr(curve)
    out = {"value": ms, "unit": "ms", "min_ms": float(np.min(wall)), "steps": steps, "warmup": warmup,
           "stages_rank0_ms": mean_stages([s[0] for s in st]),
           "e2e": {"value": ms, "unit": "ms", "h2d_bytes_per_step": int((d["nv"] + 1) * fr),
                   "d2h_bytes_per_step": int(world * pk.blob_bytes),
                   "note": "DistributedProver.prove IS the host-buffer call: witness in page-locked host memory on rank 0 -> H2D -> "
                           "NCCL broadcast -> witness map (A/B/C chains on ranks 0-2 when N >= 3) -> shard MSMs -> NCCL "
                           "all-gather of the window sums -> D2H -> host fold + assembly on rank 0; there is no separate "
                           "device-resident figure at N > 1"},
           "witness_broadcast_bytes": int((d["nv"] + 1) * fr), "allgather_bytes_per_rank": int(pk.blob_bytes),
           "chain_parallel_witness_map": bool(prover.chain_parallel),
           "sizes": {"num_constraints": d["nc"], "num_inputs": d["ni"], "num_variables": d["nv"], "domain": d["m"]},
           "gpu_launches_rank0": int(launches), "clocks": clocks, "setup_s": setup_s,
           "precomputed_tables": bool(ctx.tables), "inputs": d["source"]}
    if rank == 0:
        if fx is not None and not ctx.args.no_cpu_baseline:
            want, times, ref_st, _, _ = reference_prove(ctx.ref, fx, rs, 1, 0, 120.0)
            ok = bool(np.array_equal(want, proof))
            out["parity_checked"] = ok
            out["parity"] = "the N-rank NCCL proof equals the reference prover's proof, bit for bit, on the same inputs"
            out["cpu_baseline"] = {"value": float(np.mean(times)), "unit": "ms", "cores": ctx.ref.num_threads(),
                                   "kind": "reference", "stage_ms": ref_st,
                                   "sample": f"r1cs_gg_ppzksnark_prover<{curve}> on the same 2^{log_m} inputs, one run, no warm-up"}
            if not ok:
                raise SystemExit("bench.py: the N-rank proof differs from the reference prover's")
        elif totals is not None:
            # no reference inputs: compare with the one-GPU prover over the whole key (b200_groth16_prove) on rank 0
            full = [b.make_bases(curve, grp, n, k0=k0) for (grp, n), k0 in zip(totals, k0s)]
            one = b.Groth16ProvingKey(curve, *pts, full[0], full[1], full[2], None, full[3], full[4], cs, m)
            want, _ = one.prove(d["primary"], aux, rs[0], rs[1])
            out["matches_single_gpu"] = bool(np.array_equal(want, proof))
            out["parity_checked"] = False
            out["parity"] = "N-rank proof == one-GPU proof of the same key (b200_groth16_prove); the one-GPU prover's parity with " \
                            "the reference is what tests/test_groth16_gpu.py and the config-1 line check"
            one.free()
            del full
            if not out["matches_single_gpu"]:
                raise SystemExit("bench.py: the N-rank proof differs from the one-GPU proof")
    if fx is not None:
        fx.close()
    pk.free()
    cs.free()
    return out


# ------------------------------------------------------------------------------------------------
# G1 MSM (BASELINE configs[1]); weak scaling by base range at N > 1
# ------------------------------------------------------------------------------------------------

def bench_msm(ctx, curve, log_n, steps, warmup, wide_peak, with_plain=True, with_cold=True, ref_steps=2):
    b, dist, torch = ctx.b, ctx.dist, ctx.torch
    world, rank = ctx.world, ctx.rank
    L, cid = b.lib(), b._cid(curve)
    n = 1 << log_n
    ref = ctx.ref
    # this rank's shard of the global (world * n)-point MSM
    if ref is not None and world == 1:
        from oracle import refo
        bases = ref.make_bases(curve, refo.G1, n, 1)      # the reference's own distinct-affine-point recipe
        src = "bases from the reference (make_distinct_special), numpy scalars: the same arrays go to libff::multi_exp"
    else:
        bases = b.make_bases(curve, b.G1, n, k0=1 + rank * n)
        src = "device-generated multiples of the generator, numpy scalars"
    scalars = bench_scalars(curve, n, 1234 + rank)
    rb = b.ResidentBases(curve, b.G1, bases)
    rb_plain = b.ResidentBases(curve, b.G1, bases) if (with_plain and ctx.tables and world == 1) else None
    table_rows = rb.precompute() if ctx.tables else 0
    table_c = L.b200_bases_table_window_bits(rb.handle) if table_rows else 0
    _keep, pinned_np = ctx.pinned_copy(scalars)
    dscal = b.DeviceBuffer(scalars.nbytes)
    dscal.upload(scalars)
    W = L.b200_msm_num_windows(cid, n)
    c_plain = int(L.b200_msm_window_bits(cid, n))
    wbytes = L.b200_msm_window_bytes(cid, b.G1)
    win_t = torch.empty(W * wbytes // 8, dtype=torch.int64, device=ctx.dev)
    gather_t = torch.empty(world * W * wbytes // 8, dtype=torch.int64, device=ctx.dev) if world > 1 else None

    def finish_multi():
        dist.all_gather_into_tensor(gather_t, win_t)
        if rank == 0:
            allw = gather_t.cpu().numpy().view(np.uint64)
            return b.fold_windows(curve, b.G1, n, allw, world)
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

    for _ in range(warmup):
        r0 = step_resident()
        step_e2e()
    launches0 = b.launch_count()
    ev_ms, wall_ms, stage = ctx.timed(step_resident, steps, after=lambda r: b.last_msm_timing())
    launches = b.launch_count() - launches0
    _, e2e_wall, _ = ctx.timed(step_e2e, steps)
    r1 = step_e2e()
    if rank == 0 and not np.array_equal(r0, r1):
        raise SystemExit("bench.py: resident and e2e MSM paths disagree")
    ms_step, e2e_ms = float(np.mean(wall_ms)), float(np.mean(e2e_wall))
    total_pts = world * n
    st = mean_stages(stage)
    acc_ms = st["accumulate"]
    W_acc = table_rows if table_rows else W
    c_run = table_c if table_rows else c_plain
    mm = MM[N32_FQ[curve]]
    out = {"metric": "G1 MSM Mpts/s", "value": total_pts / (ms_step * 1e-3) / 1e6, "unit": "Mpts/s", "ms_per_step": ms_step,
           "ms_per_step_device_events": float(np.mean(ev_ms)), "scaling": "weak", "n_gpus": world,
           "higher_is_better": True, "steps": steps, "warmup": warmup,
           "config": {"workload": MSM_WORKLOAD if (curve == CURVE and log_n == 20) else f"{curve} G1 multi_exp 2^{log_n}",
                      "points_per_gpu": n, "parallelism": f"base-range shards x{world}",
                      "window_bits_run": int(c_run), "windows_run": int(W_acc), "bucket_sets_reduced": 1 if table_rows else int(W),
                      "window_bits_interface": c_plain, "windows_interface": int(W),
                      "precomputed_table_rows": int(table_rows), "inputs": src,
                      "l2": "256 MiB buffer written between timed iterations (outside the timed span)"},
           "stage_ms": st,
           "e2e": {"value": total_pts / (e2e_ms * 1e-3) / 1e6, "unit": "Mpts/s", "ms_per_step": e2e_ms,
                   "h2d_bytes_per_step": int(scalars.nbytes), "d2h_bytes_per_step": int(W * wbytes),
                   "note": "b200_msm_resident: query cached on the device, scalars from pinned host memory, window sums back "
                           "to the host, fold + normalise on the host"},
           "gpu_launches": int(launches)}
    if rank == 0:
        imads_acc = 11 * W_acc * n * mm
        ach = imads_acc / (acc_ms * 1e-3)
        whole = msm_modmuls(n, c_run, W_acc, 1 if table_rows else W) * mm
        out["roofline"] = {
            "bound": "int-pipe", "kernel": "bucket accumulation (k_accumulate + affine rounds when enabled)",
            "achieved": ach / 1e12, "peak": wide_peak / 1e12, "unit": "Tmad/s (32x32+64 multiply-adds, SURVEY.md §8d accounting: "
                                                                      "11 modmuls per mixed add)",
            "frac": ach / wide_peak, "kernel_ms": acc_ms, "algorithmic_bytes": 68.0 * W_acc * n,
            "whole_msm_achieved": whole / (ms_step * 1e-3) / 1e12, "whole_msm_frac": whole / (ms_step * 1e-3) / wide_peak}
    # cold: the literal libff::multi_exp signature — bases cross PCIe on every call (b200_msm)
    if with_cold and world == 1:
        b.multi_exp(curve, b.G1, bases, scalars)
        t = []
        for _ in range(3):
            ctx.sync_all()
            t0 = time.perf_counter()
            rc = b.multi_exp(curve, b.G1, bases, scalars)
            t.append((time.perf_counter() - t0) * 1e3)
        if not np.array_equal(rc, r0):
            raise SystemExit("bench.py: cold and resident MSM paths disagree")
        out["cold"] = {"value": n / (float(np.mean(t)) * 1e-3) / 1e6, "unit": "Mpts/s", "ms_per_step": float(np.mean(t)),
                       "h2d_bytes_per_step": int(bases.nbytes + scalars.nbytes),
                       "note": "b200_msm (multiexp.hpp:63-73 as is): bases (pageable host memory, reference layout) + scalars H2D, "
                               "base conversion, MSM without tables, result D2H — every call"}
    if rb_plain is not None:
        for _ in range(warmup):
            rb_plain.multi_exp(None, device_ptr=dscal.ptr)
        _, pw, pst = ctx.timed(lambda: rb_plain.multi_exp(None, device_ptr=dscal.ptr), steps, after=lambda r: b.last_msm_timing())
        out["without_tables"] = {"value": n / (float(np.mean(pw)) * 1e-3) / 1e6, "unit": "Mpts/s",
                                 "ms_per_step": float(np.mean(pw)), "stage_ms": mean_stages(pst),
                                 "note": f"the same MSM over the same resident key without b200_bases_precompute ({W} windows, {W} bucket sets)"}
        rb_plain.free()
    # parity + cpu baseline
    out["parity_checked"] = False
    if rank == 0 and ref is not None:
        from oracle import refo
        if world == 1:
            want, rate, rms = reference_msm_rate(ref, curve, bases, scalars, ref_steps, 1)
            ok = bool(np.array_equal(want, r0))
            out["parity"] = f"the timed table-mode 2^{log_n} MSM's point equals libff::multi_exp's on the same bases and scalars"
            out["cpu_baseline"] = {"value": rate, "unit": "Mpts/s", "cores": ref.num_threads(), "kind": "reference",
                                   "sample": f"libff::multi_exp<{curve}_G1, Fr, BDLO12_signed, special>(chunks={ref.num_threads()}) on "
                                             f"all 2^{log_n} points, {ref_steps} timed runs after 1 warm-up"}
        else:
            ok = None   # filled in by the prefix check below
        out["parity_checked"] = ok
    if world > 1:
        # parity of the N-shard fold: scalars beyond a 2^14 prefix of every shard set to zero, one extra (untimed)
        # sharded MSM, against the sum over the shards of the reference's MSM on those prefixes
        pre = 1 << 14
        z = scalars.copy()
        z[pre:] = 0
        dscal.upload(z)
        got = step_resident()
        dscal.upload(scalars)
        if rank == 0 and ref is not None:
            from oracle import refo
            acc = None
            for g_ in range(world):
                bs = b.make_bases(curve, b.G1, pre, k0=1 + g_ * n)
                sc = bench_scalars(curve, n, 1234 + g_)[:pre]
                part = ref.msm(curve, refo.G1, bs, sc, chunks=ref.num_threads())
                acc = part if acc is None else ref.group_op(curve, refo.G1, refo.GOP_ADD, acc[None, :], part[None, :])[0]
            acc = ref.group_op(curve, refo.G1, refo.GOP_TO_AFFINE, acc[None, :])[0]
            ok = bool(np.array_equal(acc, got))
            out["parity_checked"] = ok
            out["parity"] = (f"fold of the {world} shards (scalars zeroed beyond a 2^14 prefix per shard, one untimed run) == sum of "
                             f"libff::multi_exp over those prefixes")
        ctx.sync_all()
    if rank == 0 and out.get("parity_checked") is False and ref is not None:
        raise SystemExit("bench.py: MSM result differs from the reference's")
    rb.free()
    dscal.free()
    return out


# ------------------------------------------------------------------------------------------------
# NTT (BASELINE configs[2])
# ------------------------------------------------------------------------------------------------

def bench_ntt(ctx, curve, log_m, wide_peak):
    """basic_radix2_domain<Fr>::FFT at 2^log_m: resident (value), through the host-buffer C-ABI call b200_ntt (e2e),
    bit-compared with the reference's FFT of the same array, whose single run is the cpu_baseline."""
    import ctypes as C
    b, torch = ctx.b, ctx.torch
    L, cid = b.lib(), b._cid(curve)
    m = 1 << log_m
    data = bench_scalars(curve, m, 77)
    fr = data.shape[1] * 8
    buf = b.DeviceBuffer(data.nbytes)
    buf.upload(data)
    assert L.b200_ntt_dev(cid, b.FFT, C.c_void_p(buf.ptr), m, None) == 0
    got = buf.download(data.shape)
    for _ in range(3):
        assert L.b200_ntt_dev(cid, b.FFT, C.c_void_p(buf.ptr), m, None) == 0
    steps = 5
    launches0 = b.launch_count()
    b.timer_start()
    for _ in range(steps):
        assert L.b200_ntt_dev(cid, b.FFT, C.c_void_p(buf.ptr), m, None) == 0
    ms = b.timer_stop() / steps
    launches = (b.launch_count() - launches0) // steps
    buf.free()
    _keep, pn = ctx.pinned_copy(data)
    assert L.b200_ntt(cid, b.FFT, C.c_void_p(pn.ctypes.data), m, None) == 0
    e2e_ok = bool(np.array_equal(pn, got))
    e2e = []
    for _ in range(3):
        pn[:] = data
        t0 = time.perf_counter()
        assert L.b200_ntt(cid, b.FFT, C.c_void_p(pn.ctypes.data), m, None) == 0
        e2e.append((time.perf_counter() - t0) * 1e3)
    e2e_ms = float(np.mean(e2e))
    del _keep, pn
    peaks = measured_peaks()
    nc = ncu_constants().get("ntt_pass", {})
    gbs = 2 * m * fr / (ms * 1e-3) / 1e9
    mm = MM[N32_FR[curve]]
    passes = -(-log_m // 8)
    out = {"metric": "NTT elem/s", "log_m": log_m, "value": m / (ms * 1e-3), "unit": "elem/s", "ms": ms,
           "gpu_launches_per_transform": int(launches),
           "e2e": {"value": m / (e2e_ms * 1e-3), "unit": "elem/s", "ms": e2e_ms,
                   "h2d_bytes_per_step": int(data.nbytes), "d2h_bytes_per_step": int(data.nbytes),
                   "note": "b200_ntt in place on a page-locked host vector: H2D, transform, D2H inside the timed call"},
           "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
                        "frac": gbs / peaks.get("hbm_gbs"), "algorithmic_bytes": 2 * m * fr,
                        "traffic_ncu": (nc.get("dram_bytes_per_launch_at_2p24", 0) * passes * (m / float(1 << 24))) or None,
                        "traffic_source": nc.get("source"),
                        "note": "a 254-bit-field NTT is integer-pipe-bound (DESIGN.md §5): int_pipe_frac is the meaningful figure; "
                                "north_star's 60 % of HBM is unreachable on arithmetic grounds"},
           "modmul_Tmad_s": (m // 2) * log_m * mm / (ms * 1e-3) / 1e12,
           "int_pipe_frac": (m // 2) * log_m * mm / (ms * 1e-3) / wide_peak,
           "parity_checked": False}
    if ctx.ref is not None:
        want, rms = reference_ntt(ctx.ref, curve, data)
        ok = bool(np.array_equal(want, got)) and e2e_ok
        out["parity_checked"] = ok
        out["parity"] = f"all 2^{log_m} outputs (resident and host-buffer call) equal basic_radix2_domain<Fr>::FFT's on the same input"
        out["cpu_baseline"] = {"value": m / (rms * 1e-3), "unit": "elem/s", "ms": rms, "cores": ctx.ref.num_threads(),
                               "kind": "reference",
                               "sample": f"basic_radix2_domain<{curve}_Fr>::FFT (basic_radix2_domain.tcc:54-60) on the same 2^{log_m} "
                                         f"elements, one run, no warm-up (SURVEY.md §8d: run once and say so)"}
        if not ok:
            raise SystemExit("bench.py: NTT output differs from the reference's")
    return out


# ------------------------------------------------------------------------------------------------
# MSM size sweep on bls12_377 / bw6_761 G1 (BASELINE configs[4])
# ------------------------------------------------------------------------------------------------

def sweep_point(ctx, curve, log_n, ref_max_log, prefix_log=14):
    """One (curve, size) point: the 2^log_n-point MSM cut by base range over the ranks (STRONG scaling), resident
    bases with tables, scalars from the host. Parity: the whole result against the reference up to ref_max_log;
    above that a 2^prefix_log prefix per shard (other scalars zeroed, one extra untimed run)."""
    b, dist, torch = ctx.b, ctx.dist, ctx.torch
    world, rank, ref = ctx.world, ctx.rank, ctx.ref
    L, cid = b.lib(), b._cid(curve)
    n = 1 << log_n
    first, cnt = b.shard_range(n, world, rank)
    bases = b.make_bases(curve, b.G1, cnt, k0=7 + first)
    scalars = np.ascontiguousarray(bench_scalars(curve, n, log_n)[first:first + cnt]) if log_n <= 24 else \
        bench_scalars(curve, cnt, log_n * 64 + rank)
    rb = b.ResidentBases(curve, b.G1, bases)
    rows = 0
    if ctx.tables:
        try:
            rows = rb.precompute()
        except b.B200Error:
            rows = 0
    W = L.b200_msm_num_windows(cid, cnt)
    wbytes = L.b200_msm_window_bytes(cid, b.G1)
    if world > 1:
        win_t = torch.empty(W * wbytes, dtype=torch.uint8, device=ctx.dev)
        gather_t = torch.empty(world * W * wbytes, dtype=torch.uint8, device=ctx.dev)
    dscal = b.DeviceBuffer(scalars.nbytes)

    def step(sc):
        if world == 1:
            return rb.multi_exp(sc)
        rb.window_sums_dev(win_t.data_ptr(), scalars=sc)
        dist.all_gather_into_tensor(gather_t, win_t)
        if rank == 0:
            return b.fold_windows(curve, b.G1, cnt, gather_t.cpu().numpy(), world)
        torch.cuda.synchronize()
        return None

    out = step(scalars)
    _, wall, stage = ctx.timed(lambda: step(scalars), 3, after=lambda r: b.last_msm_timing())
    best = float(np.min(wall))
    st = stage[int(np.argmin(wall))]
    dscal.upload(scalars)
    res = {"curve": curve, "group": "G1", "log_n": log_n, "n_gpus": world, "scaling": "strong", "points_per_gpu": cnt,
           "e2e_ms": best, "mpts_per_s_e2e": n / best / 1e3, "stage_ms_rank0": st, "table_rows": int(rows),
           "window_bits": int(L.b200_bases_table_window_bits(rb.handle) if rows else L.b200_msm_window_bits(cid, cnt)),
           "h2d_bytes_per_step": int(scalars.nbytes), "parity_checked": False}
    if world == 1:
        r_ev, r_wall, _ = ctx.timed(lambda: rb.multi_exp(None, device_ptr=dscal.ptr), 3)
        res["resident_ms"] = float(np.min(r_wall))
        res["mpts_per_s_resident"] = n / res["resident_ms"] / 1e3
    whole = log_n <= ref_max_log
    if not whole:   # prefix check: one extra untimed run
        z = scalars.copy()
        z[min(cnt, 1 << prefix_log):] = 0
        out = step(z)
    if rank == 0 and ref is not None:
        from oracle import refo
        thr = ref.num_threads()
        if whole:
            all_bases = bases if world == 1 else b.make_bases(curve, b.G1, n, k0=7)
            all_sc = bench_scalars(curve, n, log_n)
            t0 = time.perf_counter()
            want = ref.msm(curve, refo.G1, all_bases, all_sc, chunks=thr)
            rms = (time.perf_counter() - t0) * 1e3
            res["cpu_baseline"] = {"value": n / rms / 1e3, "unit": "Mpts/s", "ms": rms, "cores": thr, "kind": "reference",
                                   "sample": f"libff::multi_exp<{curve}_G1> on all 2^{log_n} points, one run"}
            res["parity"] = "whole result == libff::multi_exp on the same bases and scalars"
        else:
            pre = 1 << prefix_log
            want = None
            t_ref = 0.0
            for g_ in range(world):
                f_, c_ = b.shard_range(n, world, g_)
                k = min(c_, pre)
                bs = b.make_bases(curve, b.G1, k, k0=7 + f_)
                sc = (bench_scalars(curve, n, log_n)[f_:f_ + k] if log_n <= 24 else bench_scalars(curve, c_, log_n * 64 + g_)[:k])
                t0 = time.perf_counter()
                part = ref.msm(curve, refo.G1, bs, sc, chunks=thr)
                t_ref += time.perf_counter() - t0
                want = part if want is None else ref.group_op(curve, refo.G1, refo.GOP_ADD, want[None, :], part[None, :])[0]
            want = ref.group_op(curve, refo.G1, refo.GOP_TO_AFFINE, want[None, :])[0]
            res["cpu_baseline"] = {"value": world * pre / t_ref / 1e6, "unit": "Mpts/s", "cores": thr, "kind": "reference",
                                   "sample": f"libff::multi_exp<{curve}_G1> on {world} x 2^{prefix_log} prefix points (the reference "
                                             f"at the full 2^{log_n} takes minutes; SURVEY.md §8d config 5)"}
            res["parity"] = (f"linearity: MSM with the scalars zeroed beyond a 2^{prefix_log} prefix of every shard == sum of the "
                             f"reference's MSMs of those prefixes")
        res["parity_checked"] = bool(np.array_equal(want, out))
        if not res["parity_checked"]:
            raise SystemExit(f"bench.py: sweep point {curve} 2^{log_n} differs from the reference")
    rb.free()
    dscal.free()
    if world > 1:
        dist.barrier()
    return res


def run_sweep(args, ctx, ref=None, reference_only=False):
    """--config 5: every (curve, size) point as one JSON line. Reference arm: the reference alone up to 2^20."""
    logs = args.sweep_logs or ([16, 18, 20, 22, 24] + ([26] if (ctx and ctx.world >= 4) else []))
    if reference_only:
        from oracle import refo
        for curve in ("bls12_377", "bw6_761"):
            for log_n in [x for x in logs if x <= 20]:
                n = 1 << log_n
                bases = ref.make_bases(curve, refo.G1, n, 7)
                sc = bench_scalars(curve, n, log_n)
                _, rate, ms = reference_msm_rate(ref, curve, bases, sc, 1, 0)
                emit({"impl": "reference", "curve": curve, "group": "G1", "log_n": log_n, "value": rate,
                      "unit": "Mpts/s", "ms": ms, "cores": ref.num_threads()})
        return
    for curve in ("bls12_377", "bw6_761"):
        for log_n in logs:
            if curve == "bw6_761" and log_n >= 26 and ctx.world < 8:
                continue
            r = sweep_point(ctx, curve, log_n, ref_max_log=20)
            if ctx.rank == 0:
                emit(r)


# ------------------------------------------------------------------------------------------------
# main
# ------------------------------------------------------------------------------------------------

def run_b200_arm(args):
    ctx = Ctx(args)
    b, world, rank = ctx.b, ctx.world, ctx.rank
    K, Wm = max(1, args.steps), max(args.warmup, 3)
    if args.config == 5:
        run_sweep(args, ctx)
        return finish(ctx)
    wide, lo, _ = b.imad_peak()
    if args.config == 4:
        curve, log_m = "bls12_377", args.config4_log_m
        steps = max(3, min(K, 5))
        if world == 1:
            pr = bench_prove_single(ctx, curve, log_m, steps, Wm, want_parity=True, ref_budget_s=1200.0)
        else:
            pr = bench_prove_sharded(ctx, curve, log_m, steps, Wm, use_reference_inputs=True)
        if rank == 0:
            pr.update({"metric": f"Groth16 prove ms @2^{log_m} {curve}", "n_gpus": world, "scaling": "strong",
                       "higher_is_better": False, "config": {"workload": config4_workload(log_m)}})
            emit(pr)
        return finish(ctx)

    # ---- config 1: the headline line ------------------------------------------------------------
    if world == 1:
        pr = bench_prove_single(ctx, CURVE, args.prove_log_m, K, Wm)
    else:
        pr = bench_prove_sharded(ctx, CURVE, args.prove_log_m, K, Wm, use_reference_inputs=True)
    line = None
    if rank == 0:
        ms = pr["value"]
        line = {"metric": METRIC if args.prove_log_m == 20 else f"Groth16 prove ms @2^{args.prove_log_m} BN254",
                "value": ms, "unit": "ms", "n_gpus": world, "steps": pr["steps"], "warmup": pr["warmup"], "ms_per_step": ms,
                "higher_is_better": False, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None,
                "dtype": "u32 limbs (254-bit modular)", "data": "synthetic",
                "config": {"workload": WORKLOAD, "parallelism": f"every query cut by base range x{world}" if world > 1 else "one GPU",
                           "inputs": pr["inputs"], "precomputed_tables": pr["precomputed_tables"],
                           "l2": "256 MiB buffer written between timed iterations (outside the timed span); the resident key "
                                 "(about 6 GiB with tables) is far larger than the 126 MB L2 anyway",
                           "scaling_note": "N > 1 shards the SAME fixed 2^20 problem: strong scaling"},
                "e2e": pr.pop("e2e"), "gpu_launches": pr.pop("gpu_launches", pr.pop("gpu_launches_rank0", 0)),
                "clocks": pr.pop("clocks"), "parity_checked": pr.get("parity_checked", False)}
        if "cpu_baseline" in pr:
            line["cpu_baseline"] = pr.pop("cpu_baseline")
        line["groth16_prove"] = pr
    msm = None
    if not args.no_msm:
        msm = bench_msm(ctx, CURVE, args.log_n, K, Wm, wide)
    if rank == 0:
        if msm is not None:
            line["msm"] = msm
            line["roofline"] = dict(msm["roofline"])
            line["roofline"]["traffic"] = None
            nc = ncu_constants().get("accumulate_alt_bn128_g1", {})
            if nc:
                # dram__bytes_read.sum + dram__bytes_write.sum of the accumulation stage's launches, from the committed
                # ncu --set full capture of this command (not measured in this run)
                line["roofline"]["traffic"] = nc.get("dram_bytes_per_launch")
                line["roofline"]["traffic_ncu"] = nc.get("dram_bytes_per_launch")
                line["roofline"]["traffic_what"] = nc.get("what")
                line["roofline"]["pipe_utilisation_ncu"] = nc.get("fmaheavy_frac")
                line["roofline"]["ncu_source"] = nc.get("source")
            line["roofline"]["peak_source"] = (
                "b200_imad_peak, measured in this run: chains of IMAD.WIDE.U32.X (the 32x32+64 multiply-add a Montgomery product "
                "is made of; it holds the FMA-heavy pipe for two issue slots, 32 per clock per SM). MEASURED_PEAKS.json has no "
                "integer figure.")
            line["roofline"]["imad32_issue_peak"] = lo / 1e12
            line["roofline"]["measured_on"] = ("the stand-alone 2^20 G1 MSM of the `msm` object (same kernel and size as the prove's A, H "
                                               "and L MSMs, which overlap inside the prove and cannot be timed one by one there)")
        if world == 1 and "algorithmic_imads" in pr:
            alg = pr["algorithmic_imads"]
            tot = alg["msm"] + alg["ntt"]
            line.setdefault("roofline", {"bound": "int-pipe", "peak": wide / 1e12})
            line["roofline"]["whole_prove"] = {
                "algorithmic_Tmad": tot / 1e12, "achieved_Tmad_s": tot / (line["value"] * 1e-3) / 1e12,
                "frac": tot / (line["value"] * 1e-3) / wide,
                "note": "five MSMs (geometry actually run, groth16_prove.msm_geometry) + seven transforms, SURVEY.md §8d "
                        "accounting, over the whole resident prove incl. host fold and assembly"}
    if world == 1 and not args.no_ntt:
        ntt = bench_ntt(ctx, CURVE, args.ntt_log_m, wide)
        line["ntt"] = ntt
    if not args.no_config4:
        steps4 = 3
        if world == 1:
            c4 = bench_prove_single(replace_ref(ctx, None), "bls12_377", args.config4_log_m, steps4, 3, want_parity=False)
        else:
            c4 = bench_prove_sharded(ctx, "bls12_377", args.config4_log_m, steps4, 3, use_reference_inputs=False)
        if rank == 0:
            c4["metric"] = f"Groth16 prove ms @2^{args.config4_log_m} bls12_377"
            c4["config"] = {"workload": config4_workload(args.config4_log_m)}
            c4["scaling"] = "strong"
            c4["n_gpus"] = world
            if world == 1 and ctx.ref is not None:
                c4.update(config4_small_parity(ctx, "bls12_377", 16))
            line["config4"] = c4
    if not args.no_sweep and world == 1:
        pts = []
        for curve in ("bls12_377", "bw6_761"):
            pts.append(sweep_point(ctx, curve, 20, ref_max_log=18, prefix_log=15))
        line["sweep"] = {"points": pts, "note": "BASELINE.json configs[4] at 2^20 per curve; the whole sweep: bench.py --config 5 "
                                                "(profiles/)"}
    if rank == 0:
        emit(line)
    finish(ctx)


def replace_ref(ctx, ref):
    """A view of ctx with another reference handle (None = synthetic device-generated inputs)."""
    import copy
    c = copy.copy(ctx)
    c.ref = ref
    return c


def config4_small_parity(ctx, curve, log_m):
    """The default line proves bls12_377 at 2^22 on a device-generated key (the reference needs minutes there: that run
    is `bench.py --config 4`, profiles/); here the same code path is checked against the reference prover at 2^log_m,
    and that reference run, scaled by nothing, is reported as the bounded cpu_baseline sample."""
    small = bench_prove_single(ctx, curve, log_m, 3, 2, want_parity=True)
    return {"parity_checked": small["parity_checked"],
            "parity": f"same prover on {curve} at 2^{log_m} against the reference prover (whole proof); the 2^22 proof itself is "
                      f"checked for determinism only in this line, and against the reference by `bench.py --config 4`",
            "cpu_baseline": dict(small.get("cpu_baseline", {}), sample_size=f"2^{log_m} of 2^22", gpu_ms_same_size=small["value"])}


def finish(ctx):
    if ctx.world > 1:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


_REAL_STDOUT = None


def emit(obj):
    """The JSON line goes to the process's real stdout; everything else that prints there (the reference's profiling lines
    '* Elements of w ...', multiexp.tcc:737-751) has been sent to stderr by main()."""
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(obj) + "\n").encode())


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)   # C-level stdout of this process (the reference library's printf) -> stderr
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
