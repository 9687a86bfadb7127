"""Per-role wait-cycle breakdown of the assessor's layers (first two blocks of every stage + stem) at P particles."""
import os, sys
# role counters exist in the generic kernel only: keep every layer on at most two epilogue groups (three run the lean kernel only)
os.environ.setdefault("PSOB_TUNING", "1"); os.environ.setdefault("PSOB_EPI3", "0")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import numpy as np, torch
import bench
from psob import engine, _ffi

P = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
net = engine.AssessorNet(bench.make_assessor().state_dict(), P)
img = torch.rand(P, 784, device="cuda")
for _ in range(2):
    net.forward(img, rescale=True)
torch.cuda.synchronize()
buf = torch.zeros(148, 32, dtype=torch.int64, device="cuda")
lib = _ffi.lib()
print("layer        jobs steps | prod: wait_empty/total | mma: wait_full wait_tmem /total | epi: wait_tfull/total | cyc/job cyc/step | epi per chunk: ld free math pub")
for op, layer in zip(net.ops, net.layers):
    if op.kind != "layer":
        continue
    pr = op.prog
    parts = pr.name.split(".")
    if len(parts) == 4 and int(parts[2]) > 1:
        continue
    src, dst = net.buf[op.src], (net.buf[op.dst] if op.dst else None)
    res = net.buf[op.res] if op.res else None
    buf.zero_()
    _ffi.check(lib.psob_layer_profile(layer.handle, buf.data_ptr()))
    _ffi.check(lib.psob_layer_run_residual(layer.handle, src.data_ptr(), src.stride(0), P, engine._ptr(dst), engine._ptr(net.logits if op.dst32 else None),
                                           engine._ptr(res), res.stride(0) if res is not None else 0, net.impl, engine._stream()), pr.name)
    torch.cuda.synchronize()
    _ffi.check(lib.psob_layer_profile(layer.handle, None))
    b = buf.cpu().numpy().astype(np.float64)
    b = b[b[:, 7] > 0]
    if not len(b):
        print(pr.name, "(no epilogue counters: direct-store epilogue)")
        continue
    m = b.mean(axis=0)
    spj = pr.n_steps / pr.n_jobs
    print(f"{pr.name:10s} {pr.n_jobs:5d} {pr.n_steps:5d} | {m[0]/m[1]:.2f} of {m[1]/1e3:8.0f}k | {m[2]/m[4]:.2f} {m[3]/m[4]:.2f} of {m[4]/1e3:8.0f}k | "
          f"{m[5]/m[6]:.2f} of {m[6]/1e3:8.0f}k | {m[4]/m[7]:8.0f} {m[4]/m[7]/spj:7.0f} | "
          f"{m[8]/m[6]:.2f} {m[9]/m[6]:.2f} {m[10]/m[6]:.2f} {m[11]/m[6]:.2f}")
