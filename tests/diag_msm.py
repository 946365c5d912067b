"""GPU diagnostic: G1 MSM over device-generated progression bases vs the discrete-log closed form,
for several (table size, prefix, copies, c) choices.  Test tooling only."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))   # tests/ -> repo root
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "contangle-zkcp_b200"))
import numpy as np
import ctg
from ctg import synthetic as S
from oracle import ark
from oracle.py import bls12_381 as B

ctx = ctg.Context(0)
a_int, b_int = 0xC0FFEE + 22, 0x5EED
rng = np.random.default_rng(7)
for (log_tab, log_n, copies, c) in [(20, 20, 1, 0), (20, 20, 2, 0), (20, 20, 0, 0), (22, 20, 0, 0), (22, 22, 0, 0), (22, 22, 1, 0), (22, 20, 4, 13), (22, 18, 0, 0)]:
    nt, n = 1 << log_tab, 1 << log_n
    t = ctx.bases_progression_g1(S.g1_gen_words(), S.fr_mont([a_int])[0], S.fr_mont([b_int])[0], nt, copies=copies, c_bits=c)
    sc = rng.integers(0, 1 << 32, size=(n, 8), dtype=np.uint64).astype(np.uint32)
    sc[:, 7] &= 0x3FFFFFFF
    t0 = time.time()
    got = ark.g1_to_py(ctx.msm(t, sc))[0]
    dt = time.time() - t0
    s_int = ark.words_to_ints(sc)
    dl = (a_int * sum(s_int) + b_int * sum(i * s for i, s in enumerate(s_int))) % B.R_MOD
    want = B.G1.mul(B.G1_GEN, dl)
    print(f"table 2^{log_tab} n 2^{log_n} copies {copies} c {c}: match={got == want} ({dt*1e3:.1f} ms)", flush=True)
    t.free()
ctx.close()
