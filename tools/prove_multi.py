#!/usr/bin/env python
"""Groth16 prove sharded over the GPUs of one box (BASELINE.json configs[3]): every query of the proving
key cut by base range, one process per GPU, window sums all-gathered over NCCL, proof assembled on rank 0.

    python tools/prove_multi.py --log-m 20                                              one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port 29555 tools/prove_multi.py --curve bls12_377 --log-m 22        N GPUs

Fixed problem, so the scaling it shows is STRONG. Inputs: the synthetic satisfiable R1CS of tools/synth_r1cs.py
(shape of generate_r1cs_example_with_field_input, num_inputs = 10, num_constraints = 2^log_m - 11) and a proving key of
distinct device-generated points of the right sizes; every rank generates only its own shard. One JSON line on rank 0.
--check also proves on rank 0 alone through b200_groth16_prove (whole key on one GPU) and compares the proofs."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def run(b, torch, dist, curve, log_m, steps, warmup, chain_parallel=None, check=False, sy=None, tables=True):
    import synth_r1cs
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    ni, m = 10, 1 << log_m
    t0 = time.time()
    if sy is None:
        sy = synth_r1cs.make(curve, m - ni - 1, ni)
    nv = sy["nv"]
    cs = b.R1CS(curve, sy["nc"], ni, nv, sy["row_ptr"], sy["col"], sy["val"])
    # the whole key would be: A [1, nv+1], B.g2, B.g1, H [.., m-1], L [.., nv-ni] as consecutive multiples of the generator
    totals = [(b.G1, nv + 1), (b.G2, nv + 1), (b.G1, nv + 1), (b.G1, m - 1), (b.G1, nv - ni)]
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
                                  world, sliced=True, B_total=nv + 1, precompute=tables)
    del shards
    prover = b.DistributedProver(pk, chain_parallel=chain_parallel)
    if rank == 0:  # the witness in page-locked memory, as bench.py's e2e contract has it
        pinned = torch.empty(sy["auxiliary"].shape, dtype=torch.int64, pin_memory=True)
        pinned.numpy().view(np.uint64)[:] = sy["auxiliary"]
        sy = dict(sy)
        sy["auxiliary"], sy["_pinned"] = pinned.numpy().view(np.uint64), pinned
    setup_s = time.time() - t0
    rs = b.random_scalars(curve, 2, seed=99)
    dev = torch.device("cuda", torch.cuda.current_device())

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    proof = None
    for _ in range(warmup):
        proof = prover.prove(sy["primary"], sy["auxiliary"], rs[0], rs[1])
    times = []
    for _ in range(steps):
        sync()
        t0 = time.perf_counter()
        p = prover.prove(sy["primary"], sy["auxiliary"], rs[0], rs[1])
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        times.append(dt)
        if rank == 0:
            assert np.array_equal(p, proof), "prover is not deterministic for fixed r, s"
    out = {"metric": f"Groth16 prove ms @2^{log_m} {curve}", "value": float(np.mean(times)), "unit": "ms",
           "min_ms": float(np.min(times)), "higher_is_better": False, "n_gpus": world, "scaling": "strong",
           "steps": steps, "warmup": warmup, "chain_parallel_witness_map": bool(prover.chain_parallel), "precomputed_tables": bool(tables),
           "sizes": {"num_constraints": sy["nc"], "num_inputs": ni, "num_variables": nv, "domain": m},
           "h2d_bytes_per_step": int((nv + 1) * b.sizeof_fr(curve)), "witness_broadcast_bytes": int((nv + 1) * b.sizeof_fr(curve)) if world > 1 else 0,
           "allgather_bytes_per_rank": int(pk.blob_bytes), "setup_s": setup_s,
           "note": "wall clock, max over ranks, around DistributedProver.prove: witness (host, rank 0) -> H2D -> NCCL broadcast -> witness map "
                   "(replicated, or the A/B/C chains on ranks 0-2 + NCCL broadcast) -> this rank's shard of the A, B(G2+G1), H, L "
                   "MSMs -> NCCL all-gather of the window sums -> host fold and proof assembly on rank 0"}
    out["stages_rank0_ms"] = {k: round(v, 3) for k, v in prover.last_stages.items()}
    if world > 1:  # every rank's split of the last step (the slowest rank sets the time)
        allst = [None] * world
        dist.all_gather_object(allst, {k: round(v, 3) for k, v in prover.last_stages.items()})
        out["stages_all_ranks_ms"] = allst
    out["note"] = out["note"].replace("(replicated, or the A/B/C chains on ranks 0-2 + NCCL broadcast)",
                                      "(replicated; or the A/B/C chains on ranks 0-2, gathered on rank 0, which hands out slices of H)")
    if check:
        ok = None
        if rank == 0:
            full = [b.make_bases(curve, grp, n, k0=k0) for (grp, n), k0 in zip(totals, k0s)]
            one = b.Groth16ProvingKey(curve, *pts, full[0], full[1], full[2], None, full[3], full[4], cs, m)
            want, _ = one.prove(sy["primary"], sy["auxiliary"], rs[0], rs[1])
            ok = bool(np.array_equal(want, proof))
            one.free()
        out["matches_single_gpu"] = ok
    pk.free()
    cs.free()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--curve", default="alt_bn128")
    ap.add_argument("--log-m", type=int, default=20)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--no-chain-parallel", action="store_true")
    ap.add_argument("--no-tables", action="store_true")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    import libsnark_b200 as b
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank), pg_options=opts)
    b.init(local_rank)
    out = run(b, torch, dist, args.curve, args.log_m, args.steps, args.warmup,
              chain_parallel=False if args.no_chain_parallel else None, check=args.check, tables=not args.no_tables)
    if (dist.get_rank() if world > 1 else 0) == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
