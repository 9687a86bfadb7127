"""Short eager run of the BASELINE config-2 step for ncu: 3 warm-up iterations + 2 measured ones.
Kernels per iteration (11): g.l1 g.l2 g.l3 d.c1 d.c2 d.c3 score_fitness pbest_minloc gbest_candidate gbest_apply pso_update."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))

import torch
import bench
from psob import engine

P = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
gen, disc = bench.make_models()
gnet = engine.GeneratorNet(gen.state_dict(), P)
dnet = engine.DiscriminatorNet(disc.state_dict(), P)
sw = engine.SwarmEngine(P, 100, generator=gnet, fitness=engine.DScoreFitness(dnet), w_inertia=0.73, w_cogn=1.496, w_soci=1.496, seed=42)
sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(42))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(iters):
    if i == iters - 2:
        e0.record()
    sw.step()
e1.record()
torch.cuda.synchronize()
print(f"P={P} iters={iters} last2_ms_per_step={e0.elapsed_time(e1) / 2:.3f} gbest={float(sw.gbest_val):.6f}")
