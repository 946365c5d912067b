"""One large MulChain proof on one GPU with a key generated on the device (no CPU leg): wall-clock median of 3.
usage: python tools/prove_large.py LOG_N"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "contangle-zkcp_b200")]
import ctg
from ctg import synthetic as S

lg = int(sys.argv[1])
ctx = ctg.Context(0)
t0 = time.perf_counter()
r1cs, zm, _ = S.mul_chain(lg)
print(f"circuit 2^{lg}: {time.perf_counter() - t0:.1f} s on the host", flush=True)
t0 = time.perf_counter()
dp = ctx.generate_parameters(r1cs, S.trapdoor()[0])
print(f"key generation: {time.perf_counter() - t0:.2f} s", flush=True)
dpk = dp.to_device_pk()
dp.free()
dr = ctx.upload_r1cs(r1cs)
zp = ctx.pinned_like(zm)
r, s = S.fr_mont([0xABCDEF])[0], S.fr_mont([0x123457])[0]
ctx.prove(dpk, dr, zp, r, s)
ts = []
for _ in range(3):
    t0 = time.perf_counter()
    _, tm = ctx.prove(dpk, dr, zp, r, s, timed=True)
    ts.append((time.perf_counter() - t0) * 1e3)
print(f"prove 2^{lg}: wall ms {sorted(ts)}  stages {tm}")
