"""Per-role wait-cycle breakdown of every pixel-GEMM layer at the BASELINE config-2 size."""
import os, sys
# role counters exist in the generic kernel only: keep every layer on at most two epilogue groups (three run the lean kernel only)
os.environ.setdefault("PSOB_TUNING", "1"); os.environ.setdefault("PSOB_EPI3", "0")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import numpy as np, torch
import bench
from psob import engine, _ffi

P = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
gen, disc = bench.make_models()
gnet = engine.GeneratorNet(gen.state_dict(), P)
dnet = engine.DiscriminatorNet(disc.state_dict(), P)
z = torch.zeros(P, gnet.in_pitch, dtype=torch.bfloat16, device="cuda")
engine.pack_latent(torch.randn(P, 100, device="cuda"), z)
for _ in range(2):
    gnet.forward(z); dnet.forward(gnet.image_padded_bf16())
torch.cuda.synchronize()
buf = torch.zeros(148, 32, dtype=torch.int64, device="cuda")
cur = z
print("layer        jobs steps | prod: wait_empty/total | mma: wait_full wait_tmem /total | epi: wait_tfull/total | cyc/item cyc/step | epi per chunk: ld free math pub")
for net, x0 in ((gnet, z), (dnet, gnet.image_padded_bf16())):
    cur = x0
    plan = dict(net.launch_plan(x0)) if hasattr(net, "launch_plan") else {}     # fused head: d.c2 runs psob_layer_run_dot, d.c3 is gone
    for layer, y, y32 in zip(net.layers, net.acts, net.acts32):
        if plan and layer.prog.name not in plan:
            continue
        buf.zero_()
        _ffi.check(_ffi.lib().psob_layer_profile(layer.handle, buf.data_ptr()))
        if plan:
            plan[layer.prog.name]()
        else:
            # the image layer as the D-score step runs it: bf16 padded image only (PROFILE_F32=1: also the fp32 image)
            layer.run(cur, P, y, y32 if (layer is not net.layers[-1] or y is None or os.environ.get("PROFILE_F32") == "1") else None)
        torch.cuda.synchronize()
        _ffi.check(_ffi.lib().psob_layer_profile(layer.handle, None))
        b = buf.cpu().numpy().astype(np.float64)
        b = b[b[:, 7] > 0]
        m = b.mean(axis=0)
        pr = layer.prog
        steps_per_item = pr.n_steps / pr.n_jobs
        print(f"{pr.name:10s} {pr.n_jobs:5d} {pr.n_steps:5d} | {m[0]/m[1]:.2f} of {m[1]/1e3:8.0f}k | {m[2]/m[4]:.2f} {m[3]/m[4]:.2f} of {m[4]/1e3:8.0f}k | "
              f"{m[5]/m[6]:.2f} of {m[6]/1e3:8.0f}k | {m[4]/m[7]:8.0f} {m[4]/m[7]/steps_per_item:7.0f} | "
              f"{m[8]/m[6]:.2f} {m[9]/m[6]:.2f} {m[10]/m[6]:.2f} {m[11]/m[6]:.2f} | warps 0-3 ld/wait/math/pub: " +
              " ".join("/".join(f"{m[16 + 4 * w + k]/m[6]:.2f}" for k in range(4)) for w in range(4)))
        cur = y
