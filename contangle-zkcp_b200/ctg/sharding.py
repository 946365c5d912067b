"""Multi-GPU plumbing for the sharded prover: one process per GPU, each holding a point-range shard of
the proving key (ctg_pk_upload_shard); a proof is `ctg_prove_partial` on every rank, one all_gather of 1152 bytes per
rank over NCCL (NVLink / NVSwitch), and `ctg_prove_combine` on rank 0.  The same code runs over gloo with CPU tensors
(tests/test_sharding_cpu.py exercises the rank arithmetic and the exchange without a GPU).

torch.distributed is plumbing here, nothing else: the payload is a byte tensor.
"""
from __future__ import annotations

import numpy as np

from . import PARTIAL_W


def shard_range(total: int, index: int, count: int):
    """[lo, lo + cnt) of `total` positions owned by shard `index` of `count` (same arithmetic as csrc/prove.cu)."""
    if count < 1 or not 0 <= index < count:
        raise ValueError("bad shard index / count")
    lo = total * index // count
    hi = total * (index + 1) // count
    return lo, hi - lo


def key_shard_ranges(num_instance: int, num_witness: int, domain_size: int, index: int, count: int):
    """Ranges a shard keeps of each extended query: ext = a/b1/b2 query ++ 3 fixed points, l, h (bit-reversed order)."""
    m = num_instance + num_witness
    return {"ext": shard_range(m + 3, index, count), "l": shard_range(num_witness, index, count),
            "h": shard_range(domain_size, index, count)}


def all_gather_partials(partial: np.ndarray, dist, device=None):
    """All-gather one (PARTIAL_W,) uint32 partial per rank -> (world, PARTIAL_W), ordered by rank.
    `dist` is torch.distributed (initialised); `device` = the rank's CUDA device for NCCL, None for gloo/CPU."""
    import torch
    partial = np.ascontiguousarray(partial, dtype=np.uint32).reshape(PARTIAL_W)
    t = torch.from_numpy(partial.view(np.int32).copy())
    if device is not None:
        t = t.to(device)
    world = dist.get_world_size()
    out = torch.empty(world * PARTIAL_W, dtype=torch.int32, device=t.device)   # flat: gloo and NCCL both accept it
    dist.all_gather_into_tensor(out, t)
    return out.cpu().numpy().view(np.uint32).reshape(world, PARTIAL_W)


def chain_owners(world: int):
    """Rank that computes each of the three witness-map chains (a, b, c): spread over the first three ranks."""
    if world < 1:
        raise ValueError("bad world size")
    return [k % world for k in range(3)]


def chain_mask(rank: int, world: int) -> int:
    """Bit k set when `rank` owns chain k (ctg_prove_partial_begin's chain_mask)."""
    return sum(1 << k for k, o in enumerate(chain_owners(world)) if o == rank)


SHARD_DEN = 1 << 20
CHAIN_COST = 0.0     # weight of one NTT chain relative to ALL five MSMs of a proof when sizing the shards.  A model of the
                     # measured chain cost (0.033: about 20 ms against 626 ms at 2^24) was tried on 8 x B200 and lost to
                     # equal shards (2^22: 37.1 vs 35.6 ms, 2^24: 108.9 vs 106.8 ms): the proof's critical path is the arrival
                     # of the chains plus the final transform and H MSM on EVERY rank, not the owners' total work.


def shard_bounds(rank: int, world: int, distribute_ntt: bool = True, chain_cost: float = CHAIN_COST):
    """(lo_num, hi_num, den) for ctg_pk_upload_range: the share of every query rank `rank` keeps.  Equal shares when the
    witness map is replicated; otherwise ranks that own NTT chains get fewer points, so that MSM share + chain work is the
    same on every rank:  share_i = (1 + 3 c) / N - k_i c  with k_i chains on rank i.  The bounds of all ranks tile [0, den)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    if not distribute_ntt or world == 1 or chain_cost <= 0:
        return rank * SHARD_DEN // world, (rank + 1) * SHARD_DEN // world, SHARD_DEN
    owners = chain_owners(world)
    k = [owners.count(i) for i in range(world)]
    share = [max((1 + 3 * chain_cost) / world - k[i] * chain_cost, 0.25 / world) for i in range(world)]
    tot = sum(share)
    cum = [0]
    for x in share:
        cum.append(cum[-1] + x / tot)
    edges = [round(c * SHARD_DEN) for c in cum]
    edges[0], edges[-1] = 0, SHARD_DEN
    return edges[rank], edges[rank + 1], SHARD_DEN


def range_of(total: int, bounds):
    """[lo, lo + cnt) of `total` positions for (lo_num, hi_num, den) -- same arithmetic as csrc/prove.cu shard_range."""
    lo_num, hi_num, den = bounds
    lo, hi = total * lo_num // den, total * hi_num // den
    return lo, hi - lo


def exchange_chains(abc, dist):
    """Broadcast chain k of `abc` (a (3, n, 8) int32 tensor: the coset evaluations of a, b, c) from its owner rank, so that
    every rank ends up with all three.  NCCL over NVLink for CUDA tensors, gloo for CPU tensors."""
    for k, owner in enumerate(chain_owners(dist.get_world_size())):
        dist.broadcast(abc[k], src=owner)
    return abc


def sharded_prove(ctx, pk_shard, r1cs, z, r, s, dist, device=None, distribute_ntt=None, phases=None):
    """ark_groth16::create_proof over `world` GPUs.  Every rank calls this with its own context and key shard and the
    same (z, r, s); rank 0 returns (A, B, C), the other ranks return None.
    distribute_ntt (default: on for NCCL with more than one rank): the three NTT chains of the witness map are computed by
    their owner ranks and broadcast over NVLink (n * 32 bytes each) instead of being replicated on every rank.
    phases: optional dict that receives this rank's wall-clock milliseconds per phase (diagnostics)."""
    import time
    world = dist.get_world_size()
    if distribute_ntt is None:
        distribute_ntt = device is not None and world > 1
    t0 = time.perf_counter()
    if distribute_ntt:
        import torch
        abc = torch.empty((3, r1cs.n, 8), dtype=torch.int32, device=device)
        ctx.prove_partial_begin(pk_shard, r1cs, z, r, s, chain_mask(dist.get_rank(), world), abc.data_ptr())
        t1 = time.perf_counter()
        exchange_chains(abc, dist)
        torch.cuda.current_stream(device).synchronize()
        t2 = time.perf_counter()
        part = ctx.prove_partial_finish(pk_shard, r1cs, r, s, abc.data_ptr())
    else:
        part = ctx.prove_partial(pk_shard, r1cs, z, r, s)
        t1 = t2 = time.perf_counter()
    t3 = time.perf_counter()
    allp = all_gather_partials(part, dist, device)
    out = ctx.prove_combine(allp, r, s) if dist.get_rank() == 0 else None
    if phases is not None:
        t4 = time.perf_counter()
        phases.update(begin_ms=(t1 - t0) * 1e3, exchange_ms=(t2 - t1) * 1e3, finish_ms=(t3 - t2) * 1e3, gather_combine_ms=(t4 - t3) * 1e3)
    return out
