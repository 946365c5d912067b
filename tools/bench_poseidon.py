"""Bulk Poseidon timing on one B200: Merkle tree over 2^k device-resident leaves and bulk two-to-one hashes, CUDA events on
the launching stream; the same tree through the host-buffer call (copies inside).  Parity is tests/test_poseidon.py's job;
the big-int checker's own speed (0.22 ms per permutation on one core) was timed there.
Usage (GPU box): python tools/bench_poseidon.py [--logs 16 20 22]"""
import argparse
import json
import os
import random
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "contangle-zkcp_b200"))
import ctg  # noqa: E402
from ctg import synthetic as S  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--logs", type=int, nargs="*", default=[16, 20, 22])
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream()
    ctx = ctg.Context(0, stream=stream.cuda_stream)
    rnd = random.Random(11)
    full_rounds, partial_rounds, alpha = 8, 31, 17        # the reference's shape (circuits-ark/src/parameters.rs:226-229)
    rounds = full_rounds + partial_rounds
    mds = S.fr_mont([rnd.randrange(S.R_MOD) for _ in range(9)]).reshape(3, 3, 8)
    rc = S.fr_mont([rnd.randrange(S.R_MOD) for _ in range(3 * rounds)]).reshape(rounds, 3, 8)
    pos = ctx.poseidon(full_rounds, partial_rounds, alpha, mds, rc)
    imad_per_s, _ = ctx.imad_peak_probe()
    sbox = bin(alpha).count("1") - 1 + alpha.bit_length() - 1
    muls = 9 * rounds + sbox * (3 * full_rounds + partial_rounds)
    out = {"params": {"full_rounds": full_rounds, "partial_rounds": partial_rounds, "alpha": alpha, "fr_products_per_permutation": muls},
           "mac32_peak_per_s": imad_per_s / 2, "trees": {}}
    for lg in args.logs:
        n = 1 << lg
        leaves = torch.randint(0, 1 << 28, (n, 8), device=dev, dtype=torch.int32)
        dig = torch.empty((n, 8), device=dev, dtype=torch.int32)
        inner = torch.empty((n, 8), device=dev, dtype=torch.int32)
        for _ in range(2):
            pos.merkle_tree_dev(leaves.data_ptr(), n, dig.data_ptr(), inner.data_ptr())
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record(stream)
        for _ in range(reps):
            pos.merkle_tree_dev(leaves.data_ptr(), n, dig.data_ptr(), inner.data_ptr())
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        perms = 2 * n - 1
        host_leaves = leaves.cpu().numpy().view(np.uint32)
        pos.merkle_tree(host_leaves)
        t0 = time.perf_counter()
        pos.merkle_tree(host_leaves)
        host_ms = (time.perf_counter() - t0) * 1e3
        out["trees"][f"2^{lg}"] = {"ms": ms, "permutations": perms, "Mperm_per_s": perms / ms / 1e3,
                                   "frac_of_mac32_peak": perms * muls * 136.0 / (ms * 1e-3) / (imad_per_s / 2),
                                   "host_buffers_ms": host_ms}
    # bulk JubJub scalar multiplication (the group part of encrypt()): 2^20 scalars on one shared base, device-resident
    d_jj = S.fr_mont([(-10240 * pow(10241, -1, S.R_MOD)) % S.R_MOD])[0]
    gen = S.fr_mont([8076246640662884909881801758704306714034609987455869804520522091855516602923,
                     13262374693698910701929044844600465831413122818447359594527400194675274060458])
    n = 1 << 20
    d_gen = torch.from_numpy(gen.view(np.int32).copy()).to(dev)
    ks = torch.randint(-(1 << 31), (1 << 31) - 1, (n, 8), device=dev, dtype=torch.int32)
    res = torch.empty((n, 2, 8), device=dev, dtype=torch.int32)
    call = lambda: ctx.te_mul_batch_dev(d_jj, d_gen.data_ptr(), 1, ks.data_ptr(), n, res.data_ptr())
    call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(3):
        call()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    out["te_mul_2^20"] = {"ms": ms, "Mscalar_mults_per_s": n / ms / 1e3}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
