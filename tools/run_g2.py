"""GPU tooling: a few G2 MSMs (for ncu captures of msm_accumulate_kernel<Fq2>)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "contangle-zkcp_b200"))
import torch
import ctg
from ctg import synthetic as S
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 18
c = int(sys.argv[2]) if len(sys.argv) > 2 else 0
dev = torch.device("cuda", 0)
ctx = ctg.Context(0, stream=torch.cuda.current_stream().cuda_stream)
n = 1 << lg
b = ctx.bases_progression_g2(S.g2_gen_words(), S.fr_mont([77 + lg])[0], S.fr_mont([3])[0], n, c_bits=c)
sc = torch.randint(0, 1 << 30, (n, 8), device=dev, dtype=torch.int32)
res = torch.zeros(96, dtype=torch.int32, device=dev)
for _ in range(3):
    ctx.msm_dev(b, sc.data_ptr(), n, res.data_ptr())
torch.cuda.synchronize()
b.free(); ctx.close()
