"""HBM-bound kernels of the swarm step at shard sizes where launch effects vanish (P = 262 144 and
P = 1 048 576): algorithmic GB/s against the measured HBM peak.  Usage: python tools/elementwise_bw.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import torch
from psob import engine

peak = 6536.0
pp = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(pp):
    peak = json.load(open(pp)).get("hbm_gbs", peak)
D = 100
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, reps=7):
    ts = []
    for _ in range(reps):
        flush.zero_()                              # evict L2 between timed launches
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


print(f"HBM peak {peak:.0f} GB/s (MEASURED_PEAKS.json); median of 7 launches, L2 flushed before each")
print(f"{'kernel':34s} {'P':>9s} {'alg. bytes/particle':>20s} {'ms':>8s} {'GB/s':>8s} {'frac':>6s}")
for P in (65536, 262144, 1048576):
    f32 = dict(dtype=torch.float32, device="cuda")
    pos, vel, pb = torch.randn(P, D, **f32), torch.randn(P, D, **f32), torch.randn(P, D, **f32)
    gb = torch.randn(D, **f32)
    z = torch.zeros(P, 128, dtype=torch.bfloat16, device="cuda")
    improved = (torch.rand(P, device="cuda") < 0.5).to(torch.uint8)
    fit, pbv = torch.rand(P, **f32) + 0.1, torch.rand(P, **f32) + 0.1
    key = torch.full((1,), 2 ** 63 - 1, dtype=torch.int64, device="cuda")
    score = torch.rand(P, **f32)
    rows = []
    rows.append(("pso_update (Philox, no pbest copy)", 5 * D * 4,
                 timed(lambda: engine.pso_update(pos, vel, pb, gb, w=0.73, c_cogn=1.496, c_soci=1.496, seed=42, iteration=3))))
    rows.append(("pso_update + latent emit + 50% pbest copy", 5 * D * 4 + 256 + 200 + 1,
                 timed(lambda: engine.pso_update(pos, vel, pb, gb, improved=improved, z=z, w=0.73, c_cogn=1.496, c_soci=1.496, seed=42, iteration=3))))
    rows.append(("pbest_minloc", 13, timed(lambda: engine.pbest_minloc(fit, pbv, improved, key))))
    rows.append(("score_fitness", 8, timed(lambda: engine.score_fitness(score, fit, "optimize_out_training"))))
    if P <= 262144:
        img, tgt = torch.rand(P, 784, **f32), torch.rand(P, 784, **f32)
        tgt16 = tgt.bfloat16()
        out = torch.empty_like(img)
        stem = torch.empty(P, 1440, dtype=torch.bfloat16, device="cuda")
        rows.append(("recon_fitness (fp32 target)", 2 * 3136 + 4, timed(lambda: engine.recon_fitness(img, tgt, fit))))
        rows.append(("recon_fitness (bf16 target)", 3136 + 1568 + 4, timed(lambda: engine.recon_fitness(img, tgt16, fit))))
        rows.append(("rescale_minmax", 2 * 3136, timed(lambda: engine.rescale_minmax(img, out))))
        from psob import _ffi
        rows.append(("pack_image_stem (+rescale)", 3136 + 2880, timed(lambda: _ffi.check(_ffi.lib().psob_pack_image_stem(
            img.data_ptr(), P, 1, stem.data_ptr(), None, engine._stream()), "pack"))))
        del img, tgt, tgt16, out, stem
    for name, nbytes, ms in rows:
        gbs = nbytes * P / ms / 1e6
        print(f"{name:34s} {P:9d} {nbytes:20d} {ms:8.4f} {gbs:8.0f} {gbs / peak:6.2f}")
