"""Probe: may the D operand of tcgen05.mma start at a TMEM column that is not a multiple of 16?
One job: three overwrite steps fill 48 columns, then accumulate steps land at columns 2, 6, 10, 18, 30."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import numpy as np, torch
from psob import engine, lowering as L, _ffi

rng = np.random.default_rng(0)
nt, n = 48, 16
offs = [0, 16, 32, 2, 6, 10, 18, 30]
tiles = [L.to_bf16_f32(rng.standard_normal((n, 64)).astype(np.float32)) for _ in offs]
steps = [(0 if i % 2 == 0 else 64, i * n, o, 0 if i < 3 else L.STEP_ACCUM) for i, o in enumerate(offs)]
jobs = [(0, len(steps), 0, 0, 0xFFFFFFFF, 0xFFFF, 0, 0)]
prog = L._finish("probe", n, nt, 1, L.ACT_NONE, 128, 0, nt, tiles, False, steps, jobs, [-1] * nt, list(range(nt)),
                 np.zeros(nt, np.float32), macs_exact=0)
P = 300
x = L.to_bf16_f32(rng.standard_normal((P, 128)).astype(np.float32))
_, ref = L.emulate(prog, x)
layer = engine.Layer(prog)
xd = torch.from_numpy(x).cuda().to(torch.bfloat16)
for impl, name in ((_ffi.IMPL_SIMT, "simt"), (_ffi.IMPL_TCGEN05, "tcgen05")):
    y = torch.zeros(P, nt, dtype=torch.float32, device="cuda")
    layer.run(xd, P, None, y, impl)
    torch.cuda.synchronize()
    err = np.abs(y.cpu().numpy() - ref).max(axis=0)
    print(name, "max |err| per column:", np.array2string(err, precision=3, max_line_width=250))
