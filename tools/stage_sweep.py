"""GPU tooling: per-stage MSM times (sort / accumulate / reduce) over sizes and window choices."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "contangle-zkcp_b200"))
import torch
import ctg
from ctg import synthetic as S

dev = torch.device("cuda", 0)
ctx = ctg.Context(0, stream=torch.cuda.current_stream().cuda_stream)
def run(group, lg, c, copies=0):
    n = 1 << lg
    mk = ctx.bases_progression_g1 if group == 1 else ctx.bases_progression_g2
    gen = S.g1_gen_words() if group == 1 else S.g2_gen_words()
    b = mk(gen, S.fr_mont([77 + lg])[0], S.fr_mont([3])[0], n, copies=copies, c_bits=c)
    sc = torch.randint(0, 1 << 30, (n, 8), device=dev, dtype=torch.int32)
    res = torch.zeros(96, dtype=torch.int32, device=dev)
    for _ in range(2):
        ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr())
    e1.record(); torch.cuda.synchronize()
    ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr(), timed=True)
    t = ctx.last_timings()
    print(f"G{group} 2^{lg} c={c:2d} copies={copies}: total {e0.elapsed_time(e1)/3:7.3f} ms  sort {t['msm_sort_ms']:.3f} acc {t['msm_accumulate_ms']:.3f} reduce {t['msm_reduce_ms']:.3f}", flush=True)
    b.free()
for lg in (14, 16, 18, 20, 22):
    for c in ([0, 8, 10, 12, 14, 16] if lg <= 16 else [0, 13, 16, 18, 20]):
        run(1, lg, c)
for lg in (14, 16, 20):
    for c in (0, 10, 13, 16, 18):
        run(2, lg, c)
ctx.close()
