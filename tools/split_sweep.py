"""Sweep ctg_ctx_set_msm_split for the headline MSM (device scalars and pinned host scalars)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "contangle-zkcp_b200")]
import numpy as np, torch, ctg
from ctg import synthetic as S
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 22
n = 1 << lg
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream()
ctx = ctg.Context(0, stream=stream.cuda_stream)
b = ctx.bases_progression_g1(S.g1_gen_words(), S.fr_mont([12345])[0], S.fr_mont([777])[0], n)
sc = torch.randint(0, 1 << 30, (n, 8), device=dev, dtype=torch.int32)
hs = torch.empty((n, 8), dtype=torch.int32, pin_memory=True); hs.copy_(sc)
hnp = hs.numpy().view(np.uint32)
res = torch.zeros(96, dtype=torch.int32, device=dev)
for parts in (1, 2, 3, 4):
    ctx.set_msm_split(parts)
    for _ in range(3):
        ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(8):
        ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr())
    e1.record(stream); torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1) / 8
    for _ in range(2):
        ctx.msm(b, hnp)
    t0 = time.perf_counter()
    for _ in range(8):
        ctx.msm(b, hnp)
    host_ms = (time.perf_counter() - t0) / 8 * 1e3
    print(f"2^{lg} parts={parts}: device scalars {dev_ms:.2f} ms   pinned host scalars {host_ms:.2f} ms", flush=True)
