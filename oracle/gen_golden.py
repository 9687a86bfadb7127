"""TEST INFRASTRUCTURE — generate `tests/golden/ref_golden.npz` from the LIVE reference.

Run in the build container (where `/root/reference` exists):

    python oracle/gen_golden.py

Every array stored here is an input to, or an output of, the reference's own classes
(`src/pso/util_pso.py`, `src/pso/util_discovery.py`, `src/utils/util_dcgan.py`,
`src/inverter/utils_ae/util_inverter.py`, `src/inverter/utils_vq_vae/*`, `src/hands_on/pso.py`)
imported unmodified through `oracle/ref_loader.py`.  The fixtures pin `oracle/port.py`
(CPU tests) and are the fixed points the CUDA path is compared with (GPU tests).
"""
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "ref_golden.npz")


def sd_to_np(prefix, sd, out):
    for k, v in sd.items():
        out[f"{prefix}/{k}"] = v.detach().cpu().numpy()


def randomize_bn(model, gen):
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            n = m.num_features
            m.running_mean.copy_(torch.randn(n, generator=gen) * 0.3)
            m.running_var.copy_(torch.rand(n, generator=gen) * 1.5 + 0.25)
            m.weight.data.copy_(torch.rand(n, generator=gen) * 1.0 + 0.5)
            m.bias.data.copy_(torch.randn(n, generator=gen) * 0.2)


class QueueDiscovery:
    """Stands in for Discovery so the real Swarm consumes a supplied fitness vector."""

    def __init__(self, values):
        self.values = list(values)
        self.k = 0

    def fitness(self, dim_space, pos):
        v = np.asarray([self.values[self.k]], dtype=np.float32)
        self.k += 1
        return v, torch.zeros(1, 1, 2, 2), None


def main():
    ref = ref_loader.load()
    U, D_, G_ = ref.util_pso, ref.util_discovery, ref.util_dcgan
    g = {}
    rng = np.random.RandomState(1234)
    tg = torch.Generator().manual_seed(1234)

    # ---- 1. teacher-forced move_particles on real Particle objects --------------------
    P, Dm = 37, 100
    pos = rng.randn(P, Dm)
    vel = (rng.randn(P, Dm) - 0.5) / 10
    pbest = pos + 0.3 * rng.randn(P, Dm)
    gbest = rng.randn(Dm)
    r1, r2 = rng.random_sample(P), rng.random_sample(P)
    w, wc, ws = 0.73, 1.496, 1.2
    new_pos, new_vel = np.empty_like(pos), np.empty_like(vel)
    orig_random = np.random.random
    for i in range(P):
        p = U.Particle(dim=Dm, pos=pos[i].copy())
        p.vel = vel[i].copy()
        p.p_best_pos = pbest[i].copy()
        draws = iter([r1[i], r2[i]])
        np.random.random = lambda: next(draws)
        try:
            p.update_velocity(w, wc, ws, gbest)
            p.update_position()
        finally:
            np.random.random = orig_random
        new_pos[i], new_vel[i] = p.pos, p.vel
    g.update({"move/pos": pos, "move/vel": vel, "move/pbest": pbest, "move/gbest": gbest,
              "move/r1": r1, "move/r2": r2, "move/w": np.array([w, wc, ws]),
              "move/new_pos": new_pos, "move/new_vel": new_vel})

    # ---- 2. pbest / gbest rules on the real Swarm with injected fitness vectors ---------
    Pq, T = 12, 4
    fit = rng.random_sample((T, Pq)).astype(np.float32) + 0.1
    fit[0, 3] = fit[0, 7] = fit[0].min()        # tie inside an iteration -> lowest index
    fit[1, 5] = fit[0].min()                     # equal to the incumbent -> never replaces
    fit[2, 9] = np.float32(0.05)                 # strict improvement
    fit[3, :] = fit[2, :]                        # nothing improves
    with tempfile.TemporaryDirectory() as td:
        qd = QueueDiscovery(fit.reshape(-1))
        sw = U.Swarm(td, qd, num_particles=Pq, n_iterations=T, dim_space=5, device="cpu")
        np.random.seed(7)
        sw.init_particles()
        pbv, gidx, gval, gpos, gdummy = [], [], [], [], []
        for t in range(T):
            sw.update_personal_best(iteration=t + 1)
            before = sw.g_best_pos
            sw.update_global_best()
            pbv.append(np.array([np.float32(p.p_best_val) for p in sw.swarm]).reshape(-1))
            idx = -1
            if sw.g_best_pos is not before:
                idx = [k for k, p in enumerate(sw.swarm) if p.p_best_pos is sw.g_best_pos][0]
            gidx.append(idx)
            gval.append(np.float32(sw.g_best_val[-1]).reshape(-1)[0])
            gpos.append(np.array(sw.g_best_pos, dtype=np.float64))
            gdummy.append(np.float32(sw.g_best_val_dummy[-1]).reshape(-1)[0])
            sw.move_particles()
        g.update({"best/fit": fit, "best/pbest_val": np.stack(pbv), "best/gbest_idx": np.array(gidx),
                  "best/gbest_val": np.array(gval), "best/gbest_val_dummy": np.array(gdummy),
                  "best/n_improvements": np.array(len(sw.g_best_val))})

    # ---- 3. models: small nets with full weights stored; three weight regimes ------------
    z_small = torch.randn(6, 10, 1, 1, generator=tg)
    for regime in ("refinit", "default", "trained"):
        torch.manual_seed({"refinit": 1, "default": 2, "trained": 3}[regime])
        gen = G_.Generator(z_dim=10, channels_img=1, features_g=16)
        if regime == "refinit":
            G_.initialize_weights(gen)
        if regime == "trained":
            randomize_bn(gen, tg)
        gen.eval()
        with torch.no_grad():
            img = gen(z_small)
        sd_to_np(f"gsmall_{regime}", gen.state_dict(), g)
        g[f"gsmall_{regime}/out"] = img.numpy()
    g["gsmall/z"] = z_small.numpy()

    torch.manual_seed(4)
    disc = G_.Discriminator(channels_img=1, features_d=16).eval()
    x_img = torch.tanh(torch.randn(6, 1, 28, 28, generator=tg))
    with torch.no_grad():
        g["dsmall/out"] = disc(x_img).numpy()
    sd_to_np("dsmall", disc.state_dict(), g)
    g["dsmall/x"] = x_img.numpy()

    torch.manual_seed(5)
    enc = ref.util_inverter.Encoder(enc_dim=10, channels_img=1, features_e=16).eval()
    with torch.no_grad():
        g["esmall/out"] = enc(x_img).numpy()
    sd_to_np("esmall", enc.state_dict(), g)

    # full-size G / D: weights regenerated from the seed by the tests (checksum stored)
    torch.manual_seed(11)
    gen_full = G_.Generator(z_dim=100, channels_img=1, features_g=64)
    randomize_bn(gen_full, torch.Generator().manual_seed(12))
    gen_full.eval()
    torch.manual_seed(13)
    disc_full = G_.Discriminator(channels_img=1, features_d=64).eval()
    z_full = torch.randn(8, 100, 1, 1, generator=torch.Generator().manual_seed(14))
    with torch.no_grad():
        img_full = gen_full(z_full)
        score_full = disc_full(img_full)
    g.update({"gfull/z": z_full.numpy(), "gfull/out": img_full.numpy(), "dfull/out": score_full.numpy(),
              "gfull/wsum": np.array(sum(float(v.double().sum()) for v in gen_full.state_dict().values())),
              "dfull/wsum": np.array(sum(float(v.double().sum()) for v in disc_full.state_dict().values()))})

    # VQ-VAE GAN decoder (config 3): the real class, its own xavier init
    torch.manual_seed(21)
    vqgan = ref.vq_model.VectorQuantizedVAE_GAN(1, 10, 32, None, features_g=16, features_d=16).eval()
    with torch.no_grad():
        g["vqdec/out"] = vqgan.decoder(z_small).numpy()
    sd_to_np("vqdec", {k: v for k, v in vqgan.state_dict().items() if k.startswith("decoder.")}, g)

    # ---- 4. Discovery.fitness / fitness_pso_inverter with a tiny assessor ---------------
    torch.manual_seed(31)
    gen_f = G_.Generator(z_dim=10, channels_img=1, features_g=16).eval()
    for n_class, tag in ((8, "multi"), (2, "binary")):
        torch.manual_seed(32 + n_class)
        cnn = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, n_class)).eval()
        classes = [0, 2, 3, 4, 6, 7, 8, 9][:n_class]
        for mode in ("optimize_out_training", "optimize_in_training"):
            disc_obj = D_.Discovery(iid_class=classes[1], model_gan=gen_f, model_cnn=cnn, device="cpu",
                                    control_pso_fitness=mode, iid_classes=classes, obj_fun_threshold=0.05)
            fits, fits_inv, resc, raw = [], [], [], []
            targets = torch.tanh(torch.randn(6, 1, 1, 28, 28, generator=tg))
            for i in range(6):
                zi = z_small[i].reshape(-1).double().numpy()      # numpy fp64 like Particle.pos
                f, img_r, img = disc_obj.fitness(10, zi)
                fi, _, _ = disc_obj.fitness_pso_inverter(10, zi, targets[i], w_ass=0.7, w_rec=1.3)
                fits.append(f); fits_inv.append(fi); resc.append(img_r.numpy()); raw.append(img.numpy())
            g[f"fit_{tag}_{mode}/fitness"] = np.concatenate(fits)
            g[f"fit_{tag}_{mode}/fitness_inv"] = np.concatenate(fits_inv)
            g[f"fit_{tag}_{mode}/targets"] = targets.numpy()
        sd_to_np(f"fit_{tag}/cnn", cnn.state_dict(), g)
        g[f"fit_{tag}/img_rescaled"] = np.concatenate(resc)
        g[f"fit_{tag}/img"] = np.concatenate(raw)
    sd_to_np("fit/gen", gen_f.state_dict(), g)

    # ---- 5. whole loop: the real Swarm.optimize, seeded numpy RNG ------------------------
    torch.manual_seed(41)
    cnn8 = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 8)).eval()
    classes = [0, 2, 3, 4, 6, 7, 8, 9]
    with tempfile.TemporaryDirectory() as td:
        disc_obj = D_.Discovery(iid_class=3, model_gan=gen_f, model_cnn=cnn8, device="cpu", iid_classes=classes)
        sw = U.Swarm(td, disc_obj, num_particles=6, n_iterations=5, dim_space=10, device="cpu",
                     tolerance=1e-5, w_inertia=0.73, w_cogn=1.496, w_soci=1.496)
        np.random.seed(99)
        hist, hp, hv, it = sw.optimize(schedule_inertia=True, early_stopping=False)
        g["loop/g_best_val"] = np.array([np.float32(v) for v in sw.g_best_val]).reshape(-1)
        g["loop/g_best_val_dummy"] = np.array([np.float32(v) for v in sw.g_best_val_dummy]).reshape(-1)
        g["loop/g_best_pos"] = np.asarray(sw.g_best_pos)
        g["loop/final_pos"] = np.stack([p.pos for p in sw.swarm])
        g["loop/final_vel"] = np.stack([p.vel for p in sw.swarm])
        g["loop/mean_mse"] = np.array(hist["mean_mse"])
        g["loop/iters"] = np.array(it)
        g["loop/hist0"] = hp["particle_0"].values
    sd_to_np("loop/cnn", cnn8.state_dict(), g)

    # ---- 6. hands_on generic-objective Swarm on the sphere function ----------------------
    def fun1(p):
        return p[0] ** 2 + p[1] ** 2 + p[2] ** 2
    np.random.seed(5)
    hs = ref.hands_on_pso.Swarm(obj_fun=fun1, num_particles=20, dim_space=3, n_iterations=100)
    hs.optimize()
    g["sphere/g_best_val"] = np.array(hs.g_best_val, dtype=np.float64)
    g["sphere/g_best_pos"] = np.asarray(hs.g_best_pos)

    # ---- 7. vq nearest codebook -----------------------------------------------------------
    xq = torch.randn(64, 10, generator=tg)
    cb = torch.randn(32, 10, generator=tg)
    g["vq/x"], g["vq/codebook"] = xq.numpy(), cb.numpy()
    g["vq/indices"] = ref.vq_function.vq(xq, cb).numpy()

    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **g)
    print(f"wrote {OUT}: {len(g)} arrays, {os.path.getsize(OUT) / 1e3:.1f} kB")


if __name__ == "__main__":
    main()
