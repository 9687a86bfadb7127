"""Per-layer CUDA-event times of the G and D stacks at the BASELINE config-2 size (knobs come from the environment)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import numpy as np, torch
import bench
from psob import engine

P = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
gen, disc = bench.make_models()
gnet = engine.GeneratorNet(gen.state_dict(), P); gnet.skip_f32 = True
dnet = engine.DiscriminatorNet(disc.state_dict(), P)
z = torch.zeros(P, gnet.in_pitch, dtype=torch.bfloat16, device="cuda")
engine.pack_latent(torch.randn(P, 100, device="cuda"), z)
for _ in range(3):
    gnet.forward(z); dnet.forward(gnet.image_padded_bf16())
torch.cuda.synchronize()
out = {}
for net, x0 in ((gnet, z), (dnet, gnet.image_padded_bf16())):
    cur = x0
    last = len(net.layers) - 1
    for i, (layer, y, y32) in enumerate(zip(net.layers, net.acts, net.acts32)):
        if i == last and net.skip_f32 and y is not None:
            y32 = None
        ts = []
        for _ in range(7):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); layer.run(cur, P, y, y32); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        out[layer.prog.name] = float(np.median(ts))
        cur = y
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    gnet.forward(z); dnet.forward(gnet.image_padded_bf16())
b.record(); torch.cuda.synchronize()
knobs = {k: v for k, v in os.environ.items() if k.startswith("PSOB_")}
print(knobs, " ".join(f"{k}={v:.3f}" for k, v in out.items()), f"| sum={sum(out.values()):.3f} fwd={a.elapsed_time(b) / 10:.3f} ms", flush=True)
