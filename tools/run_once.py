"""Small driver for ncu captures: runs W warm-up + K measured invocations of one primitive.
usage: python tools/run_once.py msm|msm2|ntt|prove LOG_N [copies] [c_bits]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "contangle-zkcp_b200")]
import numpy as np
import torch
import ctg
from ctg import synthetic as S

what, lg = sys.argv[1], int(sys.argv[2])
copies = int(sys.argv[3]) if len(sys.argv) > 3 else 1
c_bits = int(sys.argv[4]) if len(sys.argv) > 4 else 0
reps = int(os.environ.get("REPS", "2"))
dev = torch.device("cuda", 0)
ctx = ctg.Context(0, stream=torch.cuda.current_stream().cuda_stream)
n = 1 << lg
if what in ("msm", "msm2"):
    gen = S.g1_gen_words() if what == "msm" else S.g2_gen_words()
    mk = ctx.bases_progression_g1 if what == "msm" else ctx.bases_progression_g2
    b = mk(gen, S.fr_mont([12345])[0], S.fr_mont([777])[0], n, copies=copies, c_bits=c_bits)
    sc = torch.randint(0, 1 << 30, (n, 8), device=dev, dtype=torch.int32)
    res = torch.zeros(96, dtype=torch.int32, device=dev)
    for _ in range(reps):
        ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr(), timed=True)
        print(ctx.last_timings())
elif what == "ntt":
    x = torch.randint(0, 1 << 30, (n, 8), device=dev, dtype=torch.int32)
    for _ in range(reps):
        ctx.ntt_dev(x.data_ptr(), lg)
    torch.cuda.synchronize()
elif what == "prove":
    r1cs, zm, _ = S.mul_chain(lg)
    pk, _ = S.progression_pk(ctx, r1cs)
    dpk = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"], copies=copies, c_bits=c_bits)
    dr = ctx.upload_r1cs(r1cs)
    r, s = S.fr_mont([0xABCDEF])[0], S.fr_mont([0x123457])[0]
    for _ in range(reps):
        _, tm = ctx.prove(dpk, dr, zm, r, s, timed=True)
        print(tm)
elif what == "poe":            # lone PoE-shaped proofs (n_data = lg), generated key
    r1cs, zm, _ = S.poe_shape(lg)
    dp = ctx.generate_parameters(r1cs, S.trapdoor()[0])
    dpk = dp.to_device_pk()
    dp.free()
    dr = ctx.upload_r1cs(r1cs)
    r, s = S.fr_mont([0xABCDEF])[0], S.fr_mont([0x123457])[0]
    for _ in range(reps):
        _, tm = ctx.prove(dpk, dr, zm, r, s, timed=True)
        print(tm)
elif what == "batch":          # `lg` = proofs in the batch; PoE shape n_data = 1 (domain 2^13), generated key
    r1cs, zm, _ = S.poe_shape(1)
    dp = ctx.generate_parameters(r1cs, S.trapdoor()[0])
    dpk = dp.to_device_pk(c_bits=int(os.environ.get("CBITS", "0")))
    dp.free()
    dr = ctx.upload_r1cs(r1cs)
    zb = S.poe_shape_witnesses(ctx, 1, lg, seed=3)
    rs = S.fr_mont([3 + i for i in range(2 * lg)]).reshape(lg, 2, 8)
    for _ in range(reps):
        t0 = time.perf_counter()
        ctx.prove_batch(dpk, dr, zb, rs)
        print("prove_batch", lg, "proofs:", (time.perf_counter() - t0) * 1e3, "ms")
elif what == "setup":          # key generation (fixed_base_kernel, lagrange / qap kernels)
    r1cs, zm, _ = S.mul_chain(lg)
    for _ in range(reps):
        dp = ctx.generate_parameters(r1cs, S.trapdoor()[0])
        dp.free()
elif what == "te":             # bulk JubJub: shared-base table kernel, per-item kernel, hashed ElGamal
    d_jj = S.fr_mont([(-10240 * pow(10241, -1, S.R_MOD)) % S.R_MOD])[0]
    gen = S.fr_mont([8076246640662884909881801758704306714034609987455869804520522091855516602923,
                     13262374693698910701929044844600465831413122818447359594527400194675274060458])
    d_gen = torch.from_numpy(gen.view(np.int32).copy()).to(dev)
    ks = torch.randint(-(1 << 31), (1 << 31) - 1, (n, 8), device=dev, dtype=torch.int32)
    res = torch.empty((n, 2, 8), device=dev, dtype=torch.int32)
    for _ in range(reps):
        ctx.te_mul_batch_dev(d_jj, d_gen.data_ptr(), 1, ks.data_ptr(), n, res.data_ptr())          # te_mul_table_kernel
    pts = res[: n // 16].contiguous()
    for _ in range(reps):
        ctx.te_mul_batch_dev(d_jj, pts.data_ptr(), n // 16, ks.data_ptr(), n // 16, res.data_ptr())   # te_mul_kernel (a base per item)
torch.cuda.synchronize()
print("done")
