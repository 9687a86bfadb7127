"""GPU bring-up diagnostics for the tcgen05 pixel-GEMM engine: runs single layers through the
SIMT interpreter and the tcgen05 kernel and prints where (rows / columns / K slices) they differ
from the CPU emulation.  Each case runs in its own process so a device fault cannot hide the rest.

    python tools/bringup_umma.py            # all cases
    python tools/bringup_umma.py CASE       # one case in-process
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))

CASES = ["fc_k64_n16", "fc_k64_n256", "fc_k128_n256", "convt_m1", "convt_img", "conv_image_resident", "conv_k4", "conv_full", "multi_tile"]


def build_case(name):
    import numpy as np
    from psob import lowering as L
    rng = np.random.RandomState(0)
    P = 128
    if name == "fc_k64_n16":        # one step, one job, N=16
        w = rng.randn(16, 60, 1, 1).astype(np.float32) * 0.1
        prog = L.lower_conv_full("t", w, rng.randn(16).astype(np.float32), L.ACT_NONE)
    elif name == "fc_k64_n256":     # one step, N=256 (2 pixels x 128 ch), bf16 vec8 store
        w = rng.randn(64, 128, 1, 2).astype(np.float32)[:, :, :1, :] * 0.1
        w = np.ascontiguousarray(np.broadcast_to(w, (64, 128, 1, 2)))
        w = rng.randn(64, 128, 2, 2).astype(np.float32) * 0.1
        prog = L.lower_convt_from_1x1("t", w, rng.randn(128).astype(np.float32), L.ACT_RELU)
    elif name == "fc_k128_n256":    # two K-chunks accumulate
        w = rng.randn(100, 128, 7, 7).astype(np.float32) * 0.1
        prog = L.lower_convt_from_1x1("t", w, rng.randn(128).astype(np.float32), L.ACT_RELU)
    elif name == "convt_m1":
        w = rng.randn(128, 64, 4, 4).astype(np.float32) * 0.05
        prog = L.lower_convt_k4s2p1("t", w, rng.randn(64).astype(np.float32), L.ACT_RELU, 7, 7, m=1)
    elif name == "convt_img":
        w = rng.randn(64, 1, 4, 4).astype(np.float32) * 0.05
        prog = L.lower_convt_k4s2p1("t", w, rng.randn(1).astype(np.float32), L.ACT_TANH, 14, 14, m=2, image_out=True)
    elif name == "conv_image_resident":
        w = rng.randn(64, 1, 4, 4).astype(np.float32) * 0.2
        prog = L.lower_conv_image_k4s2p1("t", w, rng.randn(64).astype(np.float32), L.ACT_LEAKY)
    elif name == "conv_k4":
        w = rng.randn(128, 64, 4, 4).astype(np.float32) * 0.03
        prog = L.lower_conv_k4s2p1("t", w, rng.randn(128).astype(np.float32), L.ACT_LEAKY, 14, 14)
    elif name == "conv_full":
        w = rng.randn(1, 128, 7, 7).astype(np.float32) * 0.02
        prog = L.lower_conv_full("t", w, rng.randn(1).astype(np.float32), L.ACT_SIGMOID)
    elif name == "multi_tile":
        P = 1000
        w = rng.randn(128, 64, 4, 4).astype(np.float32) * 0.05
        prog = L.lower_convt_k4s2p1("t", w, rng.randn(64).astype(np.float32), L.ACT_RELU, 7, 7, m=1)
    else:
        raise SystemExit(f"unknown case {name}")
    x = L.to_bf16_f32(rng.randn(P, prog.in_pitch).astype(np.float32))
    return prog, x


def run_case(name):
    import numpy as np
    import torch
    from psob import engine, lowering as L, _ffi
    prog, x = build_case(name)
    y_ref, y32_ref = L.emulate(prog, x)
    xin = torch.from_numpy(x).cuda().bfloat16()
    P = x.shape[0]
    print(f"[{name}] mma_n={prog.mma_n} nt={prog.nt} cc={prog.cc} jobs={prog.n_jobs} steps={prog.n_steps} "
          f"wrows={prog.wpack.shape[0]} resident={prog.b_resident} vec8={prog.vec8} P={P}", flush=True)
    for impl, tag in ((_ffi.IMPL_SIMT, "simt"), (_ffi.IMPL_TCGEN05, "tcgen05")):
        net = engine.Network([prog], P, impl=impl)
        y, y32 = net.forward(xin)
        torch.cuda.synchronize()
        for got_t, ref, what in ((y, y_ref, "bf16"), (y32, y32_ref, "f32")):
            if ref is None:
                continue
            got = got_t.float().cpu().numpy()
            err = np.abs(got - ref)
            tol = 1e-3 + np.abs(ref) / 100
            bad = err > tol
            print(f"  {tag:8s} {what}: max_err={err.max():.4e} bad={int(bad.sum())}/{bad.size} "
                  f"ref_absmax={np.abs(ref).max():.3f} got_absmax={np.abs(got).max():.3f}", flush=True)
            if bad.any():
                rows = np.where(bad.any(axis=1))[0]
                cols = np.where(bad.any(axis=0))[0]
                print(f"    bad rows: n={len(rows)} first={rows[:12].tolist()} ; bad cols: n={len(cols)} first={cols[:16].tolist()}")
                r, c = rows[0], cols[0]
                print(f"    sample row {r}: got={got[r, c:c + 8].round(4).tolist()} ref={ref[r, c:c + 8].round(4).tolist()}")
                if np.isfinite(got).all() and np.abs(ref[bad]).min() > 1e-6:
                    ratio = got[bad] / ref[bad]
                    print(f"    got/ref on bad entries: median={np.median(ratio):.4f} min={ratio.min():.4f} max={ratio.max():.4f}")
    print(f"[{name}] done", flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        run_case(sys.argv[1])
    else:
        for c in CASES:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), c], capture_output=True, text=True, timeout=180)
            print(r.stdout[-3000:], end="")
            if r.returncode != 0:
                print(f"[{c}] exit code {r.returncode}\n{r.stderr[-1500:]}")
