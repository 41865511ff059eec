"""HBM write-only vs copy bandwidth on this GPU (context for K1's Jacobian stream, which is ~95 % writes)."""
import torch
dev = "cuda:0"
n = 1 << 30                       # 8 GiB of f64
a = torch.empty(n, dtype=torch.float64, device=dev)
b = torch.empty(n, dtype=torch.float64, device=dev)
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
ms = t(lambda: a.fill_(1.5)); print(f"fill_  8 GiB: {ms:.3f} ms  {n*8/ms/1e6:.0f} GB/s (write only)")
ms = t(lambda: a.zero_()); print(f"zero_  8 GiB: {ms:.3f} ms  {n*8/ms/1e6:.0f} GB/s (memset)")
ms = t(lambda: b.copy_(a)); print(f"copy_  8 GiB: {ms:.3f} ms  {2*n*8/ms/1e6:.0f} GB/s (read+write)")
ms = t(lambda: a.sum()); print(f"sum    8 GiB: {ms:.3f} ms  {n*8/ms/1e6:.0f} GB/s (read only)")
