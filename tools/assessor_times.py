"""Per-layer CUDA-event times of the ResNet-50 assessor on the engine (P particles), plus the torch/cuDNN
bf16 channels_last module on the same batch as a library baseline.  Usage: python tools/assessor_times.py [P]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
from oracle import port  # noqa: E402  (model definition + seeded parameters only)
from psob import engine, _ffi  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
m = port.Resnet([3, 4, 6, 3], 1, list(range(8)), 8).eval()
port.seeded_fill(m, 11)
net = engine.AssessorNet(m.state_dict(), P, shortcut=os.environ.get("PSOB_SHORTCUT"))
img = torch.rand(P, 784, device="cuda")


def timed(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(n):
        fn()
    ev[1].record()
    torch.cuda.synchronize()
    return ev[0].elapsed_time(ev[1]) / n


total = timed(lambda: net.forward(img, rescale=True))
print(f"P={P} assessor forward {total:.3f} ms  exact {net.macs_exact} MAC/particle issued {net.macs_issued}  "
      f"-> {2 * net.macs_exact * P / total / 1e9:.1f} TFLOP/s exact, {P / total * 1e3 / 1e6:.2f} M images/s")
lib = _ffi.lib()
rows = []
for op, layer in zip(net.ops, net.layers):
    if op.kind != "layer":
        continue
    pr = op.prog
    src = net.buf[op.src]
    dst = net.buf[op.dst] if op.dst else None
    res = net.buf[op.res] if op.res else None
    y32 = net.logits if op.dst32 else None

    def run():
        _ffi.check(lib.psob_layer_run_residual(layer.handle, src.data_ptr(), src.stride(0), P, engine._ptr(dst), engine._ptr(y32),
                                               engine._ptr(res), res.stride(0) if res is not None else 0, net.impl, engine._stream()), pr.name)
    t = timed(run, 3)
    rows.append((pr.name, t, pr.macs_exact, pr.macs_issued, pr.n_steps, pr.n_jobs, pr.mma_n))
for name, t, me, mi, ns, nj, mn in rows:
    print(f"{name:12s} {t:7.3f} ms  N={mn:3d} steps={ns:4d} jobs={nj:3d}  exact {2 * me * P / t / 1e9:7.1f} TFLOP/s  issued {2 * mi * P / t / 1e9:7.1f}")
print(f"sum of layers {sum(r[1] for r in rows):.3f} ms")
if os.environ.get("PSOB_TORCH_BASELINE", "1") != "0":
    mm = m.cuda().to(memory_format=torch.channels_last)
    x4 = img.view(P, 1, 28, 28)
    chunk = 8192
    with torch.no_grad():
        def lib_fwd(dtype):
            with torch.autocast("cuda", dtype=dtype):
                for s in range(0, P, chunk):
                    mm(x4[s:s + chunk])
        print(f"torch/cuDNN bf16 autocast, chunks of {chunk}: {timed(lambda: lib_fwd(torch.bfloat16), 2):.3f} ms")
        torch.backends.cudnn.allow_tf32 = True
        torch.backends.cuda.matmul.allow_tf32 = True
        def f32():
            for s in range(0, P, chunk):
                mm(x4[s:s + chunk])
        print(f"torch/cuDNN fp32 (TF32 allowed), chunks of {chunk}: {timed(f32, 2):.3f} ms")
