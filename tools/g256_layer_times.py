"""Per-layer CUDA-event times of G256 (BASELINE config 4: P = 8192, f = 64) incl. the fused reconstruction epilogue of the image
layer; A/B knobs come from the environment together with PSOB_TUNING=1.  Prints n_stages-relevant plan facts per layer too."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import numpy as np, torch
import bench
from psob import engine

P = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
gnet = engine.GeneratorNet(bench.make_g256(), P, prefix="gen")
z = torch.zeros(P, gnet.in_pitch, dtype=torch.bfloat16, device="cuda")
engine.pack_latent(torch.randn(P, 100, device="cuda"), z)
target = (torch.rand(P, 65536, device="cuda") * 2 - 1).bfloat16()
part = torch.zeros(gnet.recon_slots(), P, device="cuda")
plan = gnet.launch_plan(z, write_f32=False, recon=(target, part))
for _ in range(2):
    for _, f in plan:
        f()
torch.cuda.synchronize()
out = {}
for name, f in plan:
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); f(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    out[name] = float(np.median(ts))
knobs = {k: v for k, v in os.environ.items() if k.startswith("PSOB_") and k != "PSOB_TUNING"}
print(knobs, " ".join(f"{k.split('.')[-1]}={v:.2f}" for k, v in out.items() if not k.endswith(".ones")), f"| sum={sum(out.values()):.2f} ms", flush=True)
