"""Timing of the two small "next" rows: swarm diversity `Swarm.mse` (psob_pairwise_distance_sum) and the
VQ nearest-codebook search `vq` (psob_vq_nearest), against torch library calls on the same device."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import torch
from psob import engine


def timed(fn, n=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n


D = 100
print("Swarm.mse: mean pairwise distance, D = 100 (3 flop per pair and dimension + sqrt)")
for P in (1024, 4096, 16384, 65536):
    pos = torch.randn(P, D, device="cuda")
    t = timed(lambda: engine.mean_pairwise_distance(pos), 3)
    pairs = P * (P - 1) / 2
    line = f"  P={P:6d}: {t:9.3f} ms  {pairs / t / 1e6:8.1f} G pairs/s  {3 * D * pairs / t / 1e9:6.1f} TFLOP/s fp32"
    if P <= 16384:
        def lib():
            d = torch.cdist(pos, pos, compute_mode="donot_use_mm_for_euclid_dist")
            return float(torch.triu(d, 1).sum() / pairs)
        line += f"   | torch.cdist + triu + sum: {timed(lib, 2):9.3f} ms"
    print(line)
print("vq: nearest of K = 256 codes, D = 100 (the reference's codebook, src/inverter/utils_vq_vae/util_model.py)")
cb = torch.randn(256, D, device="cuda")
for P in (16384, 65536, 1048576):
    x = torch.randn(P, D, device="cuda")
    t = timed(lambda: engine.vq_nearest(x, cb))
    def lib():
        d = torch.addmm((cb ** 2).sum(1) + (x ** 2).sum(1, keepdim=True), x, cb.t(), alpha=-2.0)
        return d.min(1).indices
    print(f"  P={P:8d}: {t:8.3f} ms  {P / t / 1e3:8.1f} M rows/s  ({P * D * 4 / t / 1e6:6.0f} GB/s of input)   | torch addmm + min: {timed(lib):8.3f} ms")
