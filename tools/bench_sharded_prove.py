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
ap.add_argument("--replicate-ntt", action="store_true", help="every rank runs the whole witness map (no chain exchange)")
ap.add_argument("--chain-cost", type=float, default=None, help="weight of one NTT chain for the weighted shards (default: sharding.CHAIN_COST; 0 = equal shards)")
ap.add_argument("--phases", action="store_true", help="print every rank's wall-clock phase times of the last repetition")
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
dist_ntt = world > 1 and not args.replicate_ntt
for lg in args.log_n:
    r1cs, zm, _ = S.mul_chain(lg)
    t0 = time.perf_counter()
    dp = ctx.generate_parameters(r1cs, S.trapdoor()[0])   # every rank generates the same key on its GPU, keeps only its shard
    setup_ms = (time.perf_counter() - t0) * 1e3
    dpk = dp.to_device_pk(shard=sharding.shard_bounds(rank, world, dist_ntt, sharding.CHAIN_COST if args.chain_cost is None else args.chain_cost))   # owners of NTT chains keep fewer points
    dr = ctx.upload_r1cs(r1cs)
    r, s = S.fr_mont([0xABCDEF])[0], S.fr_mont([0x123457])[0]
    want = None
    if args.check and rank == 0:
        full = dp.to_device_pk()
        want = ctx.prove(full, dr, zm, r, s)
        full.free()
    dp.free()
    zm = ctx.pinned_like(zm)   # page-locked assignment, as the Rust shim packs it (ctg_host_alloc)
    for _ in range(2):
        sharding.sharded_prove(ctx, dpk, dr, zm, r, s, dist, dev if world > 1 else None, distribute_ntt=dist_ntt)
    times = []
    for _ in range(args.reps):
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        ph = {}
        got = sharding.sharded_prove(ctx, dpk, dr, zm, r, s, dist, dev if world > 1 else None, distribute_ntt=dist_ntt, phases=ph)
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], device=dev if world > 1 else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t[0]) * 1e3)
    if args.phases:
        allph = [None] * world
        dist.all_gather_object(allph, {k: round(v, 2) for k, v in ph.items()})
        if rank == 0:
            print(json.dumps({"circuit": f"MulChain 2^{lg}", "phases_ms_by_rank": allph}), flush=True)
    if rank == 0:
        times.sort()
        line = {"bench": "groth16_prove_sharded", "circuit": f"MulChain 2^{lg}", "n_gpus": world, "ms_median": times[len(times) // 2], "ms_min": times[0],
                "exchange_bytes_per_rank": 1152, "witness_map": "chains a,b,c on ranks %s, broadcast over NCCL" % sharding.chain_owners(world) if dist_ntt else "replicated", "setup_ms_rank0": setup_ms}
        if want is not None:
            line["matches_one_gpu_proof"] = bool(all(np.array_equal(g, w) for g, w in zip(got, want)))
        print(json.dumps(line), flush=True)
    ctx.pinned_free(zm)
    dpk.free(); dr.free()
ctx.close()
dist.destroy_process_group()
