"""Two warm-up forward passes and one more (G stack, then the D stack as the product runs it: fused head) at the BASELINE
config-2 size -- the ncu target: `ncu -k regex:pixel_gemm_umma --launch-skip 10 --launch-count 5 ... python tools/fwd_once.py (three passes in all)`."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import torch
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
print("layers per pass:", [p.name for p in gnet.programs] + [n for n, _ in dnet.launch_plan(gnet.image_padded_bf16())])
