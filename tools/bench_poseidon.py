"""Bulk Poseidon timing on one B200: Merkle tree over 2^k device-resident leaves and bulk two-to-one hashes, CUDA events on
the launching stream; the same tree through the host-buffer call (copies inside).  A bounded sample on the big-int checker
is timed beside it.  Usage (GPU box): python tools/bench_poseidon.py [--logs 16 20 22]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "contangle-zkcp_b200"))
import ctg  # noqa: E402
from oracle import ark  # noqa: E402  (checker / CPU sample only)
from oracle.py import poseidon as P  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--logs", type=int, nargs="*", default=[16, 20, 22])
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream()
    ctx = ctg.Context(0, stream=stream.cuda_stream)
    p = P.random_parameters(11)
    mds = ark.fr_mont_from_ints([v for row in p.mds for v in row]).reshape(3, 3, 8)
    rc = ark.fr_mont_from_ints([v for row in p.ark for v in row]).reshape(len(p.ark), 3, 8)
    pos = ctx.poseidon(p.full_rounds, p.partial_rounds, p.alpha, mds, rc)
    imad_per_s, _ = ctx.imad_peak_probe()
    sbox = bin(p.alpha).count("1") - 1 + p.alpha.bit_length() - 1
    muls = 9 * (p.full_rounds + p.partial_rounds) + sbox * (3 * p.full_rounds + p.partial_rounds)
    out = {"params": {"full_rounds": p.full_rounds, "partial_rounds": p.partial_rounds, "alpha": p.alpha, "fr_products_per_permutation": muls},
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
    # bounded CPU sample on the big-int checker (single thread)
    rows = [[i + 1, i + 2] for i in range(300)]
    t0 = time.perf_counter()
    for r in rows:
        P.crh(p, r)
    cpu = (time.perf_counter() - t0) / len(rows)
    out["cpu_bigint_checker"] = {"ms_per_permutation": cpu * 1e3, "cores": 1, "sample": "300 two-to-one hashes"}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
