"""Per-role wait-cycle breakdown of the G256 layers (BASELINE config 4: P = 8192, f = 64), image layer with the fused
reconstruction epilogue.  Knobs from the environment (with PSOB_TUNING=1)."""
import os, sys
# role counters exist in the generic kernel only: keep every layer on at most two epilogue groups (three run the lean kernel only)
os.environ.setdefault("PSOB_TUNING", "1"); os.environ.setdefault("PSOB_EPI3", "0")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import numpy as np, torch
import bench
from psob import engine, _ffi

P = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
gnet = engine.GeneratorNet(bench.make_g256(), P, prefix="gen")
z = torch.zeros(P, gnet.in_pitch, dtype=torch.bfloat16, device="cuda")
engine.pack_latent(torch.randn(P, 100, device="cuda"), z)
target = (torch.rand(P, 65536, device="cuda") * 2 - 1).bfloat16()
part = torch.zeros(gnet.recon_slots(), P, device="cuda")
plan = [(n, f) for n, f in gnet.launch_plan(z, write_f32=False, recon=(target, part)) if not n.endswith(".ones")]
for _, f in gnet.launch_plan(z, write_f32=False, recon=(target, part)):
    f()
torch.cuda.synchronize()
buf = torch.zeros(148, 32, dtype=torch.int64, device="cuda")
print("layer        jobs steps | prod: wait_empty/total | mma: wait_full wait_tmem /total | epi: wait_tfull/total | cyc/item cyc/step | epi: ld free math pub | ms")
for (name, launch), layer in zip(plan, gnet.layers):
    buf.zero_()
    _ffi.check(_ffi.lib().psob_layer_profile(layer.handle, buf.data_ptr()))
    a, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); launch(); b2.record()
    torch.cuda.synchronize()
    _ffi.check(_ffi.lib().psob_layer_profile(layer.handle, None))
    b = buf.cpu().numpy().astype(np.float64)
    b = b[b[:, 7] > 0]
    m = b.mean(axis=0)
    pr = layer.prog
    spi = pr.n_steps / pr.n_jobs
    print(f"{pr.name:10s} {pr.n_jobs:5d} {pr.n_steps:5d} | {m[0]/max(m[1],1):.2f} of {m[1]/1e3:8.0f}k | {m[2]/max(m[4],1):.2f} {m[3]/max(m[4],1):.2f} of {m[4]/1e3:8.0f}k | "
          f"{m[5]/m[6]:.2f} of {m[6]/1e3:8.0f}k | {m[6]/m[7]:8.0f} {m[6]/m[7]/spi:7.0f} | "
          f"{m[8]/m[6]:.2f} {m[9]/m[6]:.2f} {m[10]/m[6]:.2f} {m[11]/m[6]:.2f} | {a.elapsed_time(b2):.2f}", flush=True)
