"""Calibrate HBM ceilings on this box: pure write (fill), pure read (sum), copy; 4 GiB buffers, CUDA events."""
import torch
n = 1 << 30
a = torch.empty(n, dtype=torch.float32, device="cuda")
b = torch.empty(n, dtype=torch.float32, device="cuda")
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
w = t(lambda: a.fill_(1.0)); print(f"fill  (write {4*n/1e9:.1f} GB): {w:.3f} ms -> {4*n/w/1e6:.0f} GB/s")
w = t(lambda: torch.cuda.memset if False else a.zero_()); print(f"zero_ (write): {w:.3f} ms -> {4*n/w/1e6:.0f} GB/s")
r = t(lambda: a.sum()); print(f"sum   (read): {r:.3f} ms -> {4*n/r/1e6:.0f} GB/s")
c = t(lambda: b.copy_(a)); print(f"copy  (read+write {8*n/1e9:.1f} GB): {c:.3f} ms -> {8*n/c/1e6:.0f} GB/s")
