"""A few eager product iterations of BASELINE config 2 (for ncu / compute-sanitizer captures).

    python tools/step_once.py [--particles 65536] [--steps 3] [--assessor-shallow] [--recon]

Every iteration is the engine's own launch list (SwarmEngine.launch_plan): fill, 3 generator layers, fill, 2 discriminator
layers (fused head), psob_fitness_tail, psob_gbest_local, psob_update = 10 launches.  --assessor-shallow adds one forward of a
shallow instance of the reference's bottleneck ResNet (layers [1,1,1,1]); --recon one config-3 style iteration (reconstruction
epilogue + tail).  Prints the global best so that a run under a tool can be compared with a plain run."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import torch

from psob import engine
from src.utils import util_dcgan as M
from src.pso import util_cnn

ap = argparse.ArgumentParser()
ap.add_argument("--particles", type=int, default=65536)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--assessor-shallow", action="store_true")
ap.add_argument("--recon", action="store_true")
a = ap.parse_args()
P, D = a.particles, 100
torch.manual_seed(42)
gen, disc = M.Generator(D, 1, 64), M.Discriminator(1, 64)
M.initialize_weights(gen)
M.initialize_weights(disc)
gnet = engine.GeneratorNet(gen.eval().state_dict(), P)
dnet = engine.DiscriminatorNet(disc.eval().state_dict(), P)
sw = engine.SwarmEngine(P, D, generator=gnet, fitness=engine.DScoreFitness(dnet), w_inertia=0.73, w_cogn=1.496, w_soci=1.496, seed=42,
                        need_image_f32=False)
sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(42))
for _ in range(a.steps):
    sw.step()
torch.cuda.synchronize()
print(f"P={P}: {a.steps} iterations, launches per iteration {len(sw.launch_plan())}, gbest {float(sw.gbest_val):.7f}")
if a.recon:
    tgt = torch.rand(P, 784, device="cuda") * 2 - 1
    g2 = engine.GeneratorNet(gen.state_dict(), P)
    s2 = engine.SwarmEngine(P, D, generator=g2, fitness=engine.ReconFitness(tgt), seed=1, need_image_f32=False)
    s2.init_particles(generator=torch.Generator(device="cuda").manual_seed(3))
    s2.step()
    torch.cuda.synchronize()
    print(f"recon iteration: gbest {float(s2.gbest_val):.7f}")
if a.assessor_shallow:
    cnn = util_cnn.Resnet(util_cnn.resnet_block, [1, 1, 1, 1], 1, list(range(8)), 8).eval()
    net = engine.AssessorNet(cnn.state_dict(), P)
    img = torch.rand(P, 784, device="cuda")
    logits = net.forward(img, rescale=True)
    torch.cuda.synchronize()
    print(f"shallow assessor: logits {tuple(logits.shape)}, finite {bool(torch.isfinite(logits).all())}")
