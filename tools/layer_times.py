"""Per-launch CUDA-event times of the G and D stacks at the BASELINE config-2 size (A/B knobs come from the environment
together with PSOB_TUNING=1).  Usage: [PSOB_TUNING=1 PSOB_<KNOB>=v] python tools/layer_times.py [P]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import numpy as np, torch
import bench
from psob import engine

P = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
gen, disc = bench.make_models()
gnet = engine.GeneratorNet(gen.state_dict(), P)
dnet = engine.DiscriminatorNet(disc.state_dict(), P)
z = torch.zeros(P, gnet.in_pitch, dtype=torch.bfloat16, device="cuda")
engine.pack_latent(torch.randn(P, 100, device="cuda"), z)
plan = gnet.launch_plan(z, write_f32=False) + dnet.launch_plan(gnet.image_padded_bf16(), finish=False)
for _ in range(3):
    for _, f in plan:
        f()
torch.cuda.synchronize()
out = {}
for name, f in plan:
    ts = []
    for _ in range(7):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); f(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    out[name] = float(np.median(ts))
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    for _, f in plan:
        f()
b.record(); torch.cuda.synchronize()
knobs = {k: v for k, v in os.environ.items() if k.startswith("PSOB_")}
print(knobs, " ".join(f"{k}={v:.3f}" for k, v in out.items() if not k.endswith(".ones")), f"| sum={sum(out.values()):.3f} fwd={a.elapsed_time(b) / 10:.3f} ms", flush=True)
