"""GPU (-m gpu): whole iterations of the swarm engine and the drop-in Swarm / Discovery classes
against the oracle, teacher-forced stage by stage (free-running trajectories are chaotic)."""
import math

import numpy as np
import pytest
import torch

from oracle import port
from conftest import load_sd

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from psob import engine
    engine.require_cuda()
    return engine


def _models(seed=0, trained=True):
    torch.manual_seed(seed)
    gen = port.Generator(100, 1, 64)
    disc = port.Discriminator(1, 64)
    if trained:
        port.randomize_bn_stats(gen, torch.Generator().manual_seed(seed + 1))
    return gen.eval(), disc.eval()


def test_iteration_dscore_teacher_forced(engine):
    P, D = 512, 100
    gen, disc = _models()
    gnet = engine.GeneratorNet(gen.state_dict(), P)
    dnet = engine.DiscriminatorNet(disc.state_dict(), P)
    sw = engine.SwarmEngine(P, D, generator=gnet, fitness=engine.DScoreFitness(dnet), w_inertia=0.73, w_cogn=1.496,
                            w_soci=1.496, seed=42)
    sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(1))
    pos0, vel0 = sw.pos.cpu().numpy().copy(), sw.vel.cpu().numpy().copy()
    sw.step()
    torch.cuda.synchronize()
    # stage 1: fitness vs the oracle (fp32 torch on CPU) — bf16 generator/discriminator tolerance
    fit_ref, img_ref = port.fitness_dscore(gen, disc, torch.from_numpy(pos0), D)
    img = gnet.image().cpu().numpy()
    assert np.abs(img - img_ref.reshape(P, 784).numpy()).max() <= 2e-2
    np.testing.assert_allclose(sw.fit.cpu().numpy(), fit_ref, atol=5e-3, rtol=0)
    # stage 2: personal / global best from the engine's OWN fitness vector — bit-exact indices and values
    fit = sw.fit.cpu().numpy()
    pbv_ref, improved = port.pbest_update(fit, np.full(P, np.inf, np.float32))
    gval, gidx = port.gbest_scan(pbv_ref, np.inf)
    assert np.array_equal(sw.pbest_val.cpu().numpy(), pbv_ref)
    assert float(sw.gbest_val) == gval
    assert np.array_equal(sw.gbest_pos.cpu().numpy(), pos0[gidx])
    # stage 3: the move with in-kernel Philox, iteration counter = 1 — bit-exact vs the fp32 restatement
    r1, r2 = port.philox_r1r2(42, 1, 0, P)
    p_ref, v_ref = port.pso_move(pos0, vel0, pos0, pos0[gidx], r1, r2, 0.73, 1.496, 1.496, dtype=np.float32)
    assert np.array_equal(sw.pos.cpu().numpy(), p_ref) and np.array_equal(sw.vel.cpu().numpy(), v_ref)
    assert sw.iteration_state()[0] == 1


def test_graph_replay_equals_eager(engine):
    P, D = 1024, 100
    gen, disc = _models(3)
    def make():
        gnet = engine.GeneratorNet(gen.state_dict(), P)
        dnet = engine.DiscriminatorNet(disc.state_dict(), P)
        sw = engine.SwarmEngine(P, D, generator=gnet, fitness=engine.DScoreFitness(dnet), w_inertia=0.73, w_cogn=1.496,
                                w_soci=1.496, seed=7)
        sw.inertia_decay = 0.99
        sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(2))
        return sw
    a, b = make(), make()
    a.run(6, use_graph=False)
    b.run(6, use_graph=True)
    torch.cuda.synchronize()
    assert torch.equal(a.pos, b.pos) and torch.equal(a.vel, b.vel) and torch.equal(a.pbest_val, b.pbest_val)
    assert torch.equal(a.gbest_pos, b.gbest_pos) and a.iteration_state() == b.iteration_state()
    ha, hb = a.history(), b.history()
    assert np.array_equal(ha[0], hb[0]) and len(ha[0]) == 6
    assert abs(a.iteration_state()[1] - 0.73 * 0.99 ** 5) < 1e-6          # decay applied from the 2nd iteration on
    assert np.all(np.diff(ha[0]) <= 0)                                    # global best never gets worse


def test_recon_fitness_vqvae_config(engine, golden):
    """BASELINE config 3 shape: VQ-VAE-GAN decoder, per-particle targets, fitness = MSE + eps."""
    P, D = 256, 100
    torch.manual_seed(5)
    dec = port.VQDecoderGAN(1, 100, 64).eval()
    tgt = torch.rand(P, 784, generator=torch.Generator().manual_seed(6)) * 2 - 1
    gnet = engine.GeneratorNet(dec.state_dict(), P, prefix="decoder")
    sw = engine.SwarmEngine(P, D, generator=gnet, fitness=engine.ReconFitness(tgt.cuda(), w_rec=1.0, eps=0.1), seed=1)
    sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(3))
    pos0 = sw.pos.cpu().numpy().copy()
    sw.step()
    fit_ref, _ = port.fitness_inverter(dec, torch.from_numpy(pos0), tgt.view(P, 1, 28, 28), D, w_ass=0.0, w_rec=1.0)
    np.testing.assert_allclose(sw.fit.cpu().numpy(), fit_ref, atol=5e-3, rtol=0)


def test_dropin_swarm_discovery_api(engine, golden, tmp_path):
    from src.pso import util_pso, util_discovery
    from src.utils import util_dcgan
    gen = util_dcgan.Generator(10, 1, 16)
    gen.load_state_dict(load_sd(golden, "fit/gen"))
    gen = gen.cuda().eval()
    cnn = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 8))
    cnn.load_state_dict(load_sd(golden, "fit_multi/cnn"))
    cnn = cnn.cuda().eval()
    classes = [0, 2, 3, 4, 6, 7, 8, 9]
    # single-particle API against the live-reference fitness values (D = 10 is not a multiple of 4: scalar pack path)
    for mode in ("optimize_out_training", "optimize_in_training"):
        disc = util_discovery.Discovery(iid_class=2, model_gan=gen, model_cnn=cnn, device="cuda:0", control_pso_fitness=mode,
                                        iid_classes=classes, obj_fun_threshold=0.05)
        z = golden["gsmall/z"].reshape(6, 10).astype(np.float64)
        for i in range(6):
            f, img_r, img = disc.fitness(10, z[i])
            assert f.shape == (1,) and f.dtype == np.float32 and img.shape == (1, 1, 28, 28)
            assert abs(float(f[0]) - float(golden[f"fit_multi_{mode}/fitness"][i])) <= 2e-2
            tgt = torch.from_numpy(golden[f"fit_multi_{mode}/targets"][i])
            fi, _, _ = disc.fitness_pso_inverter(10, z[i], tgt, w_ass=0.7, w_rec=1.3)
            assert abs(float(fi[0]) - float(golden[f"fit_multi_{mode}/fitness_inv"][i])) <= 3e-2
        assert float(img_r.min()) == 0.0 and float(img_r.max()) == 1.0
    with pytest.raises(ValueError):
        util_discovery.Discovery(2, gen, cnn, "cuda:0", control_pso_fitness="bogus", iid_classes=classes).fitness(10, z[0])
    # whole loop through the reference-shaped Swarm (dim_space 10: padded to 12 inside the engine)
    gen12 = gen
    disc = util_discovery.Discovery(iid_class=3, model_gan=gen12, model_cnn=cnn, device="cuda:0", iid_classes=classes)
    sw = util_pso.Swarm(str(tmp_path), disc, num_particles=32, n_iterations=5, dim_space=10, device="cuda:0",
                        tolerance=1e-5, w_inertia=0.73, w_cogn=1.496, w_soci=1.496)
    history, hp, hv, it = sw.optimize(schedule_inertia=True, early_stopping=False)
    assert it == 6 and len(history["mean_mse"]) == 5 and len(sw.g_best_val_dummy) == 5
    assert sorted(hp.keys())[0] == "particle_0" and hp["particle_0"].shape == (6, 10) and hv["particle_31"].shape == (6, 10)
    assert all(np.diff([float(v[0]) for v in sw.g_best_val]) < 0) and sw.g_best_pos.shape == (10,)
    assert abs(sw.w_inertia - 0.73 * 0.99 ** 4) < 1e-6
    # diversity metric equals the reference's O(P^2) python loop
    pos = np.stack([p.pos for p in sw.swarm])
    assert abs(sw.mse() - port.mean_pairwise_distance(pos)) <= 1e-6 * sw.mse()
    assert sw.swarm[3].p_best_val.shape == (1,) and len(sw.swarm[0].history) == 6
    # discriminator passed as the assessor -> discriminator-score fitness (extension)
    d = util_dcgan.Discriminator(1, 16)
    d.load_state_dict(load_sd(golden, "dsmall"))
    d = d.cuda().eval()
    disc = util_discovery.Discovery(iid_class=None, model_gan=gen12, model_cnn=d, device="cuda:0")
    fit, _, img = disc.fitness_batch(torch.randn(64, 10, device="cuda"))
    pd_ = port.Discriminator(1, 16).eval()
    pd_.load_state_dict(load_sd(golden, "dsmall"))
    with torch.no_grad():
        ref = port.score_to_fitness(pd_(img.cpu()).reshape(-1), "optimize_out_training").numpy()
    np.testing.assert_allclose(fit.cpu().numpy(), ref, atol=5e-3, rtol=0)


def test_models_are_dropins_for_reference_checkpoints(engine, golden):
    from src.utils import util_dcgan
    import src.models as models
    assert models.Generator is util_dcgan.Generator
    g = util_dcgan.Generator(10, 1, 16)
    ref_keys = [k[len("gsmall_trained/"):] for k in golden if k.startswith("gsmall_trained/gen.")]
    assert sorted(g.state_dict().keys()) == sorted(ref_keys)              # same keys as the reference module
    g.load_state_dict(load_sd(golden, "gsmall_trained"))
    out = g.cuda()(torch.from_numpy(golden["gsmall/z"]).cuda())
    assert out.shape == (6, 1, 28, 28)
    assert np.abs(out.cpu().numpy() - golden["gsmall_trained/out"]).max() <= 2e-2
    d = util_dcgan.Discriminator(1, 16)
    d.load_state_dict(load_sd(golden, "dsmall"))
    s = d.cuda()(torch.from_numpy(golden["dsmall/x"]).cuda())
    assert s.shape == (6, 1, 1, 1)
    np.testing.assert_allclose(s.cpu().numpy(), golden["dsmall/out"], atol=5e-3, rtol=0)
    e = util_dcgan.Encoder(10, 1, 16)
    e.load_state_dict(load_sd(golden, "esmall"))
    assert e.cuda()(torch.from_numpy(golden["dsmall/x"]).cuda()).shape == (6, 10, 1, 1)


def test_sphere_known_answer_hands_on_swarm(engine):
    """src/hands_on/example_pso.py fun1: x^2+y^2+z^2 -> global best goes to 0."""
    from src.hands_on import pso
    s = pso.Swarm(obj_fun=lambda p: p[0] ** 2 + p[1] ** 2 + p[2] ** 2, num_particles=20, dim_space=3, n_iterations=100)
    s.optimize()
    assert s.g_best_val[-1] < 1e-2 and np.all(np.diff(s.g_best_val) < 0) and s.g_best_pos.shape == (3,)
    sb = pso.Swarm(obj_fun=lambda p: (p ** 2).sum(dim=1), num_particles=4096, dim_space=3, n_iterations=60, batched=True, tolerance=0.0)
    sb.optimize()
    assert sb.g_best_val[-1] < 1e-4


def test_inverter_swarm_and_on_disk_history_format(engine, golden, tmp_path):
    """`Swarm.optimize(ood_patient=..., encoder=...)` (src/pso/util_pso.py:93-112,167-193): particles start at E(img), the
    objective is the inverter loss; then the history pickle the PSO scripts write (src/training/pso_inverter.py:342-343)
    read back the way `upload_pso_particles` does (src/pso/util_pso_analysis.py:16-34)."""
    import pickle
    from src.pso import util_pso, util_discovery
    from src.utils import util_dcgan
    from src.inverter.utils_ae import util_inverter
    gen = util_dcgan.Generator(10, 1, 16)
    gen.load_state_dict(load_sd(golden, "fit/gen"))
    enc = util_inverter.Encoder(10, 1, 16)
    enc.load_state_dict(load_sd(golden, "esmall"))
    cnn = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 8))
    cnn.load_state_dict(load_sd(golden, "fit_multi/cnn"))
    gen, enc, cnn = gen.cuda().eval(), enc.cuda().eval(), cnn.cuda().eval()
    tg = torch.Generator().manual_seed(21)
    ood_patient = [(torch.tanh(torch.randn(1, 1, 28, 28, generator=tg)), torch.tensor([0])) for _ in range(12)]
    classes = [0, 2, 3, 4, 6, 7, 8, 9]
    disc = util_discovery.Discovery(iid_class=3, model_gan=gen, model_cnn=cnn, device="cuda:0", iid_classes=classes)
    sw = util_pso.Swarm(str(tmp_path), disc, num_particles=8, n_iterations=3, dim_space=10, device="cuda:0",
                        tolerance=1e-5, w_inertia=0.73, w_cogn=1.496, w_soci=1.496)
    history, hp, hv, it = sw.optimize(schedule_inertia=False, early_stopping=False, ood_patient=ood_patient, encoder=enc)
    assert it == 4 and len(hp) == 8 and hp["particle_7"].shape == (4, 10)
    # row 0 of every particle's history is the encoder's latent of its target image (fp32 oracle, bf16 engine)
    pe = port.Encoder(10, 1, 16).eval()
    pe.load_state_dict(load_sd(golden, "esmall"))
    tgt = torch.cat([im for im, _ in ood_patient[:8]])
    with torch.no_grad():
        z0 = pe(tgt).reshape(8, 10).numpy()
    got0 = np.stack([hp[f"particle_{i}"].values[0] for i in range(8)])
    assert np.abs(got0 - z0).max() <= 2e-2 * max(1.0, np.abs(z0).max())
    # the first global best is the minimum of the inverter loss at those positions (oracle evaluated at the engine's z0)
    pg = port.Generator(10, 1, 16).eval()
    pg.load_state_dict(load_sd(golden, "fit/gen"))
    pc = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 8)).eval()
    pc.load_state_dict(load_sd(golden, "fit_multi/cnn"))
    ass = np.concatenate([port.fitness_assessor(pg, pc, torch.from_numpy(got0[i]).float(), 10, class_index=2)[0] for i in range(8)])
    fit_ref, _ = port.fitness_inverter(pg, torch.from_numpy(got0).float(), tgt, 10, assessor_fitness=ass)
    assert abs(float(np.asarray(sw.g_best_val_dummy[0]).reshape(-1)[0]) - float(fit_ref.min())) <= 3e-2
    # on-disk format: dict particle_k -> DataFrame[(iterations + 1) x D]; the consumer takes `.iloc[iteration, :]`
    path = tmp_path / "particles_position_ood_class_0.pkl"
    with open(path, "wb") as f:
        pickle.dump(hp, f)
    with open(path, "rb") as f:
        back = pickle.load(f)
    data = np.ones((8, 10), dtype="float32")
    for particle_idx, p_key in enumerate(back.keys()):
        data[particle_idx, :] = back[p_key].iloc[-1, :]
    final = np.stack([p.pos for p in sw.swarm]).astype(np.float32)
    assert list(back.keys()) == [f"particle_{i}" for i in range(8)] and np.array_equal(data, final)


def test_full_size_step_size_independent_properties(engine):
    """BASELINE configs[1] at its full size (65 536 particles, Generator(100,1,64) + Discriminator(1,64)), where the
    oracle cannot run: properties that do not depend on the size.  (1) A particle's fitness does not depend on which
    other particles share its batch: the swarm evaluated whole, as two shards with global indices, and in reversed
    order gives the same bits per particle.  (2) pbest / gbest are the strict-< rules applied to that vector.
    (3) The Philox move depends on the GLOBAL index only.  (4) 48 particles spot-checked against the oracle."""
    P, D = 65536, 100
    gen, disc = _models(5)
    g = torch.Generator(device="cuda").manual_seed(11)
    pos0 = torch.randn(P, D, generator=g, device="cuda")
    vel0 = (torch.randn(P, D, generator=g, device="cuda") - 0.5) / 10

    def one_step(pos, vel, offset):
        n = pos.shape[0]
        gnet = engine.GeneratorNet(gen.state_dict(), n)
        dnet = engine.DiscriminatorNet(disc.state_dict(), n)
        sw = engine.SwarmEngine(n, D, generator=gnet, fitness=engine.DScoreFitness(dnet), w_inertia=0.73, w_cogn=1.496,
                                w_soci=1.496, seed=42, index_offset=offset)
        sw.init_particles(pos=pos, vel=vel)
        sw.evaluate()
        fit = sw.fit.clone()
        sw.personal_and_global_best()
        return sw, fit

    whole, fit = one_step(pos0, vel0, 0)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(fit).all()) and 0.1 <= float(fit.min()) and float(fit.max()) <= 1.1
    # (1) batch-composition independence, bit for bit
    h = P // 2
    _, fit_a = one_step(pos0[:h].contiguous(), vel0[:h].contiguous(), 0)
    _, fit_b = one_step(pos0[h:].contiguous(), vel0[h:].contiguous(), h)
    assert torch.equal(torch.cat([fit_a, fit_b]), fit)
    _, fit_r = one_step(pos0.flip(0).contiguous(), vel0.flip(0).contiguous(), 0)
    assert torch.equal(fit_r.flip(0), fit)
    # (2) the reference's rules on the engine's own fitness vector (src/pso/util_pso.py:122-124,135-151)
    f = fit.cpu().numpy()
    gval, gidx = port.gbest_scan(f, np.inf)
    assert np.array_equal(whole.pbest_val.cpu().numpy(), f) and float(whole.gbest_val) == gval
    assert gidx == int(np.flatnonzero(f == f.min())[0])
    assert torch.equal(whole.gbest_pos[:D], pos0[gidx])
    # (3) the move: shards with global indices reproduce the whole swarm's trajectory bit for bit
    whole.move()
    r1, r2 = port.philox_r1r2(42, 1, 0, P)
    rows = np.r_[0:8, h - 4:h + 4, P - 8:P]
    p_ref, v_ref = port.pso_move(pos0.cpu().numpy()[rows], vel0.cpu().numpy()[rows], pos0.cpu().numpy()[rows],
                                 pos0[gidx].cpu().numpy(), r1[rows], r2[rows], 0.73, 1.496, 1.496, dtype=np.float32)
    assert np.array_equal(whole.pos.cpu().numpy()[rows, :D], p_ref) and np.array_equal(whole.vel.cpu().numpy()[rows, :D], v_ref)
    # (4) spot check against the fp32 oracle
    idx = torch.randperm(P, generator=torch.Generator().manual_seed(1))[:48]
    fit_ref, _ = port.fitness_dscore(gen, disc, pos0[idx.cuda()].cpu(), D)
    np.testing.assert_allclose(f[idx.numpy()], fit_ref, atol=5e-3, rtol=0)
