"""GPU tooling: G1 MSM stage times at 2^20 / 2^22 / 2^24 for the library named by CTG_LIB (compare builds of the accumulate kernel)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "contangle-zkcp_b200"))
import torch
import ctg
from ctg import synthetic as S
dev = torch.device("cuda", 0)
ctx = ctg.Context(0, stream=torch.cuda.current_stream().cuda_stream)
print("library:", ctg.LIB_PATH)
for lg in (20, 22, 24):
    n = 1 << lg
    b = ctx.bases_progression_g1(S.g1_gen_words(), S.fr_mont([77 + lg])[0], S.fr_mont([3])[0], n)
    sc = torch.randint(0, 1 << 30, (n, 8), device=dev, dtype=torch.int32)
    res = torch.zeros(96, dtype=torch.int32, device=dev)
    for _ in range(3):
        ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr())
    e1.record(); torch.cuda.synchronize()
    ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr(), timed=True)
    t = ctx.last_timings()
    print(f"G1 2^{lg}: total {e0.elapsed_time(e1)/5:7.3f} ms  sort {t['msm_sort_ms']:.3f} acc {t['msm_accumulate_ms']:.3f} reduce {t['msm_reduce_ms']:.3f}", flush=True)
    b.free()
ctx.close()
