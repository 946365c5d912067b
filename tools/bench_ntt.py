"""NTT timing table (CUDA events on the launching stream): forward transform at 2^10 .. 2^24, and the witness-map chain
(ifft -> coset_fft on three vectors, coset_ifft on one) through a MulChain prove's timings.  CTG_NTT_LEGACY=1 selects the
round-1 pass kernel for an A/B comparison.   usage: python tools/bench_ntt.py [sizes...]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "contangle-zkcp_b200")]
import torch
import ctg

sizes = [int(a) for a in sys.argv[1:]] or [10, 12, 14, 16, 18, 20, 22, 24]
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream()
ctx = ctg.Context(0, stream=stream.cuda_stream)
imad, _ = ctx.imad_peak_probe()
out = {"legacy": os.environ.get("CTG_NTT_LEGACY", "0") == "1", "mac32_peak_T": imad / 2e12, "rows": {}}
for lg in sizes:
    x = torch.randint(0, 1 << 30, (1 << lg, 8), device=dev, dtype=torch.int32)
    for inverse, coset in ((0, 0), (1, 1)):
        for _ in range(3):
            ctx.ntt_dev(x.data_ptr(), lg, inverse=bool(inverse), coset=bool(coset))
        torch.cuda.synchronize()
        reps = 20 if lg <= 20 else 8
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            ctx.ntt_dev(x.data_ptr(), lg, inverse=bool(inverse), coset=bool(coset))
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        macs = (1 << lg) / 2 * lg * 136.0
        out["rows"][f"2^{lg}" + ("_coset_ifft" if inverse else "_fft")] = {"ms": round(ms, 4), "GBps_algorithmic": round(64.0 * (1 << lg) / ms / 1e6, 1),
                                                                          "frac_mac32": round(macs / (ms * 1e-3) / (imad / 2), 3)}
    del x
print(json.dumps(out))
