"""Random-gather bandwidth of HBM on this GPU: rows of 32/64/96/128/192 bytes from a multi-GB table (torch.index_select).
Sizing evidence for the batched-affine level-1 kernel."""
import torch, time
dev = torch.device("cuda", 0)
for row_bytes in (32, 64, 96, 128, 192, 256):
    w = row_bytes // 4
    rows = (4 << 30) // row_bytes
    tab = torch.empty((rows, w), dtype=torch.float32, device=dev).normal_()
    k = 1 << 25
    idx = torch.randint(0, rows, (k,), device=dev)
    out = torch.empty((k, w), dtype=torch.float32, device=dev)
    for _ in range(2):
        torch.index_select(tab, 0, idx, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        torch.index_select(tab, 0, idx, out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"row {row_bytes:4d} B: {k / ms / 1e6:8.2f} G rows/s  useful read {k * row_bytes / ms / 1e6:8.1f} GB/s (+ same written sequentially)")
    del tab, out, idx
