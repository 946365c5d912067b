"""world_size-2 gloo test of the multi-GPU plumbing (contangle-zkcp_b200/ctg/sharding.py) on CPU.

The GPU work of a rank (ctg_prove_partial over its point range) is stood in for by the oracle evaluating the SAME
decomposition: extended queries a ++ [alpha, delta, O] etc., bit-reversed h with slot n-1 = -rs on delta, ranges from
sharding.key_shard_ranges.  What is under test is the host logic: range arithmetic, payload layout, the all_gather and
its rank order, and that summing the shards' partials and assembling gives exactly the oracle's create_proof."""
import os
import random
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _bitrev(j, log_n):
    return int(format(j, f"0{log_n}b")[::-1], 2) if log_n else 0


def _xyzz_from_affine(ark, aff, words):
    """affine (words,) -> XYZZ (2 * words,): ZZ = ZZZ = 1 (Montgomery), infinity -> zeros"""
    out = np.zeros(2 * words, dtype=np.uint32)
    if aff.any():
        out[:words] = aff
        one = ark.fq_to_mont(ark.ints_to_words([1], 12))[0]
        half = words // 2
        out[words:words + 12] = one           # ZZ  (c0 for G2)
        out[words + half:words + half + 12] = one  # ZZZ (c0 for G2)
    return out


def _affine_from_xyzz(x, words):
    return x[:words].copy()   # ZZ = ZZZ = 1 by construction in this test


def _worker(rank, world, port, q):
    import sys
    for p in (ROOT, os.path.join(ROOT, "contangle-zkcp_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ctg import sharding, PARTIAL_W
        from ctg import synthetic as S
        from oracle import ark
        from oracle.py import bls12_381 as B
        from oracle.py import groth16 as G
        from tests import util
        rnd = random.Random(777)        # same stream on every rank
        log_n = 5
        cs, z = G.mul_chain(log_n, rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD))
        td = G.Trapdoor(*[rnd.randrange(1, B.R_MOD) for _ in range(5)], B.G1.mul(B.G1_GEN, 11), B.G2.mul(B.G2_GEN, 13))
        pk = util.trapdoor_pk_arrays(cs, td)
        r1 = ark.r1cs_from_py(cs)
        r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
        zm = ark.fr_mont_from_ints(z)
        ni, nw, n = cs.num_instance, cs.num_vars - cs.num_instance, cs.domain_size
        m = ni + nw
        rg = sharding.key_shard_ranges(ni, nw, n, rank, world)
        # the decomposition csrc/prove.cu uses
        O1, O2 = np.zeros(24, np.uint32), np.zeros(48, np.uint32)
        a_ext = np.vstack([pk["a_query"], pk["alpha_g1"], pk["delta_g1"], O1])
        b1_ext = np.vstack([pk["b_g1_query"], pk["beta_g1"], O1, pk["delta_g1"]])
        b2_ext = np.vstack([pk["b_g2_query"], pk["beta_g2"], O2, pk["delta_g2"]])
        z_ext = ark.ints_to_words(list(z) + [1, r, s], 8)
        h = ark.fr_ints_from_mont(ark.witness_map(r1, zm))
        hbr = [h[_bitrev(j, log_n)] for j in range(n)]
        hbr[n - 1] = (-r * s) % B.R_MOD
        h_ext = np.vstack([pk["h_query"][_bitrev(j, log_n)] if _bitrev(j, log_n) < n - 1 else pk["delta_g1"] for j in range(n)])
        hbr_w = ark.ints_to_words(hbr, 8)
        (elo, ecnt), (llo, lcnt), (hlo, hcnt) = rg["ext"], rg["l"], rg["h"]
        pa = ark.msm_g1(a_ext[elo:elo + ecnt], z_ext[elo:elo + ecnt])
        pb1 = ark.msm_g1(b1_ext[elo:elo + ecnt], z_ext[elo:elo + ecnt])
        pb2 = ark.msm_g2(b2_ext[elo:elo + ecnt], z_ext[elo:elo + ecnt])
        pl = ark.msm_g1(pk["l_query"][llo:llo + lcnt], z_ext[ni + llo:ni + llo + lcnt]) if lcnt else O1
        ph = ark.msm_g1(h_ext[hlo:hlo + hcnt], hbr_w[hlo:hlo + hcnt])
        part = np.concatenate([_xyzz_from_affine(ark, p, 24) for p in (pa, pb1, pl, ph)] + [_xyzz_from_affine(ark, pb2, 48)])
        assert part.shape == (PARTIAL_W,)
        allp = sharding.all_gather_partials(part, dist)
        assert allp.shape == (world, PARTIAL_W)
        assert np.array_equal(allp[rank], part)          # rank order preserved
        # distributed witness map: each rank computes the chains it owns (oracle stand-in for ctg_prove_partial_begin),
        # the exchange fills in the rest, and the finish step reproduces the oracle's h on every rank
        import torch
        mask = sharding.chain_mask(rank, world)
        assert sum(sharding.chain_mask(k, world) for k in range(world)) == 7
        abc = torch.zeros((3, n, 8), dtype=torch.int32)
        zi = list(z)
        for k, rows in enumerate((cs.a, cs.b, cs.c)):
            if not mask & (1 << k):
                continue
            ev = [sum(cf * zi[j] for cf, j in row) % B.R_MOD for row in rows] + [0] * (n - len(rows))
            if k == 0:
                for j in range(ni):
                    ev[cs.num_constraints + j] = zi[j]
            v = ark.ntt_fr(ark.ntt_fr(ark.fr_mont_from_ints(ev), inverse=True), coset=True)
            abc[k] = torch.from_numpy(v.view(np.int32).copy())
        sharding.exchange_chains(abc, dist)
        ev = [ark.fr_ints_from_mont(abc[k].numpy().view(np.uint32)) for k in range(3)]
        zinv = pow((pow(7, n, B.R_MOD) - 1) % B.R_MOD, -1, B.R_MOD)
        hq = [(x * y - w) % B.R_MOD * zinv % B.R_MOD for x, y, w in zip(*ev)]
        h2 = ark.fr_ints_from_mont(ark.ntt_fr(ark.fr_mont_from_ints(hq), inverse=True, coset=True))
        assert h2 == h, "distributed witness map differs from the oracle's"
        if rank == 0:
            def total(off, words, add):
                acc = np.zeros(words, np.uint32)
                for k in range(world):
                    p = _affine_from_xyzz(allp[k][off:off + 2 * words], words)
                    acc = p if not acc.any() else (acc if not p.any() else add(acc, p))
                return acc
            A, B1, L, H = (total(48 * i, 24, ark.g1_add) for i in range(4))
            B2 = total(192, 48, ark.g2_add)
            sw, rw = ark.ints_to_words([s], 8)[0], ark.ints_to_words([r], 8)[0]
            C = ark.g1_add(ark.g1_add(ark.g1_mul(A, sw), ark.g1_mul(B1, rw)), ark.g1_add(L, H))
            want = ark.create_proof(pk, r1, zm, ark.fr_mont_from_ints([r])[0], ark.fr_mont_from_ints([s])[0])
            ok = np.array_equal(A, want[0]) and np.array_equal(B2, want[1]) and np.array_equal(C, want[2])
            q.put(("ok" if ok else "proof mismatch", rank))
        else:
            q.put(("ok", rank))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((f"rank {rank}: {e}\n{traceback.format_exc()}", rank))
    finally:
        dist.destroy_process_group()


def test_shard_ranges_cover_everything():
    import ctg.sharding as sh
    for total in (0, 1, 7, 64, 1000003):
        for count in (1, 2, 3, 8):
            pos = 0
            for k in range(count):
                lo, cnt = sh.shard_range(total, k, count)
                assert lo == pos
                pos += cnt
            assert pos == total
    with pytest.raises(ValueError):
        sh.shard_range(10, 2, 2)


def test_weighted_shard_bounds_tile_everything():
    import ctg.sharding as sh
    for world in (1, 2, 3, 4, 8):
        for dist_ntt in (False, True):
            b = [sh.shard_bounds(r, world, dist_ntt, 0.033 if dist_ntt else 0.0) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == sh.SHARD_DEN and all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            for total in (0, 1, 7, 1 << 22, (1 << 24) + 3):
                pos = 0
                for r in range(world):
                    lo, cnt = sh.range_of(total, b[r])
                    assert lo == pos
                    pos += cnt
                assert pos == total
    # owners of NTT chains keep fewer points; the equal split is reproduced when the witness map is replicated
    b8 = [sh.shard_bounds(r, 8, True, 0.033) for r in range(8)]
    sizes = [hi - lo for lo, hi, _ in b8]
    assert sizes[0] == sizes[1] == sizes[2] < sizes[3] and abs(sizes[3] - sizes[7]) <= 1
    assert sh.shard_bounds(1, 4, distribute_ntt=False) == (sh.SHARD_DEN // 4, sh.SHARD_DEN // 2, sh.SHARD_DEN)
    assert sh.range_of(1000, sh.shard_bounds(1, 4, False)) == sh.shard_range(1000, 1, 4)
    assert sh.shard_bounds(2, 8) == (2 * sh.SHARD_DEN // 8, 3 * sh.SHARD_DEN // 8, sh.SHARD_DEN)   # default: equal shards


def test_chain_owners():
    import ctg.sharding as sh
    assert sh.chain_owners(1) == [0, 0, 0] and sh.chain_owners(2) == [0, 1, 0] and sh.chain_owners(8) == [0, 1, 2]
    assert [sh.chain_mask(r, 2) for r in range(2)] == [0b101, 0b010]
    assert [sh.chain_mask(r, 8) for r in range(8)] == [1, 2, 4, 0, 0, 0, 0, 0]


def test_two_rank_sharded_proof_over_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(rk, 2, port, q)) for rk in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(res[0] == "ok" for res in results), results
