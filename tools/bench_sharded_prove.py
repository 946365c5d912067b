#!/usr/bin/env python
"""Sharded Groth16 prove over N GPUs (BASELINE.json configs[3]): one rank per GPU under torchrun, point-range key shards,
one all_gather of 1152 B per rank, rank 0 finalises.  Prints one JSON line per size on rank 0.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/bench_sharded_prove.py --log-n 20 22
With WORLD_SIZE unset it runs the same code on one GPU (shard 0 of 1)."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "contangle-zkcp_b200")]
import numpy as np
import torch
import ctg
from ctg import synthetic as S, sharding

ap = argparse.ArgumentParser()
ap.add_argument("--log-n", type=int, nargs="+", default=[20])
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--check", action="store_true", help="compare with the one-GPU proof of rank 0 (needs the full key on rank 0)")
args = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
import torch.distributed as dist
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
else:
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:29533", rank=0, world_size=1)
ctx = ctg.Context(local, stream=torch.cuda.current_stream().cuda_stream)
for lg in args.log_n:
    r1cs, zm, _ = S.mul_chain(lg)
    pk, _ = S.progression_pk(ctx, r1cs)          # every rank builds the same synthetic key, keeps only its shard on the device
    dpk = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"], shard=(rank, world))
    dr = ctx.upload_r1cs(r1cs)
    r, s = S.fr_mont([0xABCDEF])[0], S.fr_mont([0x123457])[0]
    want = None
    if args.check and rank == 0:
        full = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
        want = ctx.prove(full, dr, zm, r, s)
        full.free()
    del pk
    for _ in range(2):
        sharding.sharded_prove(ctx, dpk, dr, zm, r, s, dist, dev if world > 1 else None)
    times = []
    for _ in range(args.reps):
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        got = sharding.sharded_prove(ctx, dpk, dr, zm, r, s, dist, dev if world > 1 else None)
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], device=dev if world > 1 else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t[0]) * 1e3)
    if rank == 0:
        times.sort()
        line = {"bench": "groth16_prove_sharded", "circuit": f"MulChain 2^{lg}", "n_gpus": world, "ms_median": times[len(times) // 2], "ms_min": times[0],
                "exchange_bytes_per_rank": 1152}
        if want is not None:
            line["matches_one_gpu_proof"] = bool(all(np.array_equal(g, w) for g, w in zip(got, want)))
        print(json.dumps(line), flush=True)
    dpk.free(); dr.free()
ctx.close()
dist.destroy_process_group()
