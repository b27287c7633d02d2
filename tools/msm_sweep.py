"""BASELINE.json configs[4]: MSM size sweep on bls12_377 and bw6_761 G1 (plus alt_bn128), resident bases, at 1/2/4/8
GPUs, with the reference's multi_exp timed on the host cores up to --ref-max-log (parity checked where both run).

    python tools/msm_sweep.py [--curves bls12_377 bw6_761] [--min-log 16] [--max-log 24] [--ref-max-log 20] [--tables]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29571 \
           tools/msm_sweep.py ...                                                      N GPUs

Under torchrun the 2^log_n points are cut by base range over the ranks (b200_shard_range, STRONG scaling): every rank
generates and uploads only its shard, computes its window sums from host scalars, the ranks all-gather them over NCCL
and rank 0 folds; the time is the max over ranks of the wall clock around that. One JSON object per (curve, size)."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libsnark_b200 as b  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--curves", nargs="+", default=["bls12_377", "bw6_761"])
    ap.add_argument("--min-log", type=int, default=16)
    ap.add_argument("--max-log", type=int, default=24)
    ap.add_argument("--ref-max-log", type=int, default=20)
    ap.add_argument("--tables", action="store_true", help="b200_bases_precompute where the table fits")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch = dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    b.init(local_rank)
    ref = None
    if rank == 0:
        try:
            from oracle import refo
            ref = refo.RefOracle()
            ref.set_num_threads(len(os.sched_getaffinity(0)))
        except (OSError, FileNotFoundError):
            pass
    for curve in args.curves:
        cid = b._cid(curve)
        for log_n in range(args.min_log, args.max_log + 1, 2):
            n = 1 << log_n
            first, cnt = b.shard_range(n, world, rank)
            bases = b.make_bases(curve, b.G1, cnt, k0=7 + first)
            small = log_n <= args.ref_max_log
            if small:   # identical global scalar vector on every rank, so that the reference can check the whole MSM
                all_scalars = b.random_scalars(curve, n, seed=log_n)
                scalars = np.ascontiguousarray(all_scalars[first:first + cnt])
            else:
                scalars = b.random_scalars(curve, cnt, seed=log_n * 64 + rank)
            rb = b.ResidentBases(curve, b.G1, bases)
            rows = 0
            if args.tables:
                try:
                    rows = rb.precompute()
                except b.B200Error:
                    rows = 0
            W = b.lib().b200_msm_num_windows(cid, cnt)
            wbytes = b.lib().b200_msm_window_bytes(cid, b.G1)
            if world > 1:
                win_t = torch.empty(W * wbytes, dtype=torch.uint8, device="cuda")
                gather_t = torch.empty(world * W * wbytes, dtype=torch.uint8, device="cuda")

            def step():
                if world == 1:
                    return rb.multi_exp(scalars)
                rb.window_sums_dev(win_t.data_ptr(), scalars=scalars)
                dist.all_gather_into_tensor(gather_t, win_t)
                if rank == 0:
                    return b.fold_windows(curve, b.G1, cnt, gather_t.cpu().numpy(), world)
                torch.cuda.synchronize()
                return None

            out = step()
            best, stage = 1e30, None
            for _ in range(3):
                if world > 1:
                    torch.cuda.synchronize()
                    dist.barrier()
                t0 = time.perf_counter()
                out = step()
                dt = (time.perf_counter() - t0) * 1e3
                if world > 1:
                    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    dt = float(t.item())
                if dt < best:
                    best, stage = dt, b.last_msm_timing()
            line = {"curve": curve, "group": "G1", "log_n": log_n, "n_gpus": world, "scaling": "strong",
                    "points_per_gpu": cnt, "e2e_ms": best, "device_ms_rank0": stage["total"], "stage_ms_rank0": stage,
                    "mpts_per_s_e2e": n / best / 1e3, "table_rows": rows,
                    "window_bits": int(b.lib().b200_msm_window_bits(cid, cnt))}
            if world == 1:
                line["mpts_per_s_resident"] = n / stage["total"] / 1e3
            if ref is not None and small:
                all_bases = bases if world == 1 else b.make_bases(curve, b.G1, n, k0=7)
                t0 = time.perf_counter()
                want = ref.msm(curve, 1, all_bases, all_scalars, chunks=ref.num_threads())
                line["reference_ms"] = (time.perf_counter() - t0) * 1e3
                line["reference_threads"] = ref.num_threads()
                line["bit_exact"] = bool(np.array_equal(out, want))
                line["speedup_e2e"] = line["reference_ms"] / best
            if rank == 0:
                print(json.dumps(line), flush=True)
            rb.free()
            del bases, scalars
            if world > 1:
                dist.barrier()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
