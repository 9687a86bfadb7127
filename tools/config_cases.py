"""BASELINE configs 3 and 4 at their full sizes on one B200 (parity-test cases, not bench lines): whole PSO steps with the
reconstruction fitness, CUDA-event timed, plus a spot check of a few particles against the fp32 oracle.
  config 3: VQ-VAE-GAN decoder, P = 16 384, per-particle fp32 targets [P,1,28,28], w_ass = 0
  config 4: G256(z=100, f=64),   P = 8 192,  per-particle bf16 targets [P,1,256,256], w_ass = 0
Usage: python tools/config_cases.py [3|4|all]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import numpy as np, torch
from oracle import port
from psob import engine

which = sys.argv[1] if len(sys.argv) > 1 else "all"
D = 100


def run(name, model, prefix, P, n_pix, target_dtype, steps=10):
    torch.manual_seed(42)
    gnet = engine.GeneratorNet(model.state_dict(), P, prefix=prefix)
    g = torch.Generator(device="cuda").manual_seed(42)
    target = (torch.rand(P, n_pix, generator=g, device="cuda") * 2 - 1).to(target_dtype)
    sw = engine.SwarmEngine(P, D, generator=gnet, fitness=engine.ReconFitness(target, w_rec=1.0, eps=0.1), w_inertia=0.73,
                            w_cogn=1.496, w_soci=1.496, seed=42)
    sw.init_particles(generator=g)
    pos0 = sw.pos[:, :D].clone()
    sw.evaluate()
    torch.cuda.synchronize()
    idx = torch.randperm(P, generator=torch.Generator().manual_seed(0))[:4]
    fit_ref, img_ref = port.fitness_inverter(model, pos0[idx.cuda()].cpu(), target[idx.cuda()].float().cpu().view(4, 1, *([int(n_pix ** 0.5)] * 2)), D)
    fit = sw.fit[idx.cuda()].cpu().numpy()
    img = gnet.image()[idx.cuda()].cpu().numpy()
    print(f"{name}: spot check of 4 particles: max|dfit| = {np.abs(fit - fit_ref).max():.2e}, max|dimg| = "
          f"{np.abs(img - img_ref.reshape(4, -1).numpy()).max():.2e}")
    sw.run(3, use_graph=False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); sw.run(steps, use_graph=False); b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    flops = 2.0 * gnet.macs_exact * P
    print(f"{name}: P = {P}: {ms:.3f} ms/step, {P / ms * 1e3 / 1e6:.3f} M particle-evals/s, generator {flops / ms / 1e9:.1f} TFLOP/s on exact MACs "
          f"({gnet.macs_exact} MAC/particle), gbest {float(sw.gbest_val):.5f}")
    ts = []
    cur = sw.z
    for layer, y, y32 in zip(gnet.layers, gnet.acts, gnet.acts32):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); layer.run(cur, P, y, y32); e1.record()
        torch.cuda.synchronize()
        ts.append((layer.prog.name, e0.elapsed_time(e1), 2.0 * layer.prog.macs_exact * P))
        cur = y
    print("   " + "  ".join(f"{n} {t:.3f} ms ({f / t / 1e9:.0f} TF/s)" for n, t, f in ts))


if which in ("3", "all"):
    torch.manual_seed(3)
    run("config 3 (VQ-VAE-GAN decoder)", port.VQDecoderGAN(1, 100, 64).eval(), "decoder", 16384, 784, torch.float32)
if which in ("4", "all"):
    torch.manual_seed(4)
    m = port.G256(100, 1, 64).eval()
    port.initialize_weights(m)
    port.randomize_bn_stats(m, torch.Generator().manual_seed(5))
    run("config 4 (G256)", m, "gen", 8192, 65536, torch.bfloat16, steps=3)
