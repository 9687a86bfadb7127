"""GPU (-m gpu): BASELINE configs 2, 3 and 4 at their FULL sizes through the product iteration (fused fitness tail,
reconstruction loss in the image epilogue), where the oracle cannot run the whole swarm: size-independent properties
(batch-composition independence bit for bit, the strict-< rules on the engine's own fitness vector) plus a spot check of
48 particles against the fp32 oracle (image: max-abs <= 2e-2 AND rel-L2 <= 5e-3; fitness against port.fitness_inverter).
Measured errors are recorded next to every gate."""
import numpy as np
import pytest
import torch

from oracle import port

pytestmark = pytest.mark.gpu

D = 100


@pytest.fixture(scope="module")
def engine():
    from psob import engine
    engine.require_cuda()
    return engine


def _recon_step(engine, model, prefix, pos, target, offset=0, need_image=False):
    n = pos.shape[0]
    gnet = engine.GeneratorNet(model.state_dict(), n, prefix=prefix)
    sw = engine.SwarmEngine(n, D, generator=gnet, fitness=engine.ReconFitness(target, w_rec=1.0, eps=0.1), w_inertia=0.73,
                            w_cogn=1.496, w_soci=1.496, seed=42, index_offset=offset, need_image_f32=need_image)
    sw.init_particles(pos=pos, vel=torch.zeros_like(pos))
    sw.evaluate_and_best()
    torch.cuda.synchronize()
    return sw


def _check_recon_config(engine, model, prefix, P, n_pix, tdtype, n_spot, img_gate, fit_gate):
    g = torch.Generator(device="cuda").manual_seed(42)
    pos = torch.randn(P, D, generator=g, device="cuda")
    target = (torch.rand(P, n_pix, generator=g, device="cuda") * 2 - 1).to(tdtype)
    whole = _recon_step(engine, model, prefix, pos, target, need_image=True)
    fit = whole.fit.clone()
    assert bool(torch.isfinite(fit).all()) and float(fit.min()) >= 0.1
    # (1) batch-composition independence, bit for bit: two shards with global indices, and the reversed swarm
    h = P // 2
    a = _recon_step(engine, model, prefix, pos[:h].contiguous(), target[:h].contiguous(), 0)
    b = _recon_step(engine, model, prefix, pos[h:].contiguous(), target[h:].contiguous(), h)
    assert not a.write_f32                                   # the product form never materialises the fp32 image
    assert torch.equal(torch.cat([a.fit, b.fit]), fit)
    del a, b
    r = _recon_step(engine, model, prefix, pos.flip(0).contiguous(), target.flip(0).contiguous(), 0)
    assert torch.equal(r.fit.flip(0), fit)
    del r
    # (2) the reference's rules on the engine's own fitness vector (src/pso/util_pso.py:122-124,135-151)
    f = fit.cpu().numpy()
    gval, gidx = port.gbest_scan(f, np.inf)
    assert np.array_equal(whole.pbest_val.cpu().numpy(), f) and float(whole.gbest_val) == gval
    assert gidx == int(np.flatnonzero(f == f.min())[0]) and torch.equal(whole.gbest_pos[:D], pos[gidx])
    # (3) spot check against the fp32 oracle (Discovery.fitness_pso_inverter with w_ass = 0, src/pso/util_discovery.py:84-101)
    idx = torch.randperm(P, generator=torch.Generator().manual_seed(1))[:n_spot]
    side = int(round(n_pix ** 0.5))
    fit_ref, img_ref = port.fitness_inverter(model, pos[idx.cuda()].cpu(), target[idx.cuda()].float().cpu().view(n_spot, 1, side, side), D)
    img = whole.gen.image()[idx.cuda()].cpu().numpy()
    ref = img_ref.reshape(n_spot, -1).numpy()
    img_err = float(np.abs(img - ref).max())
    rel = float(np.linalg.norm(img - ref) / np.linalg.norm(ref))
    fit_err = float(np.abs(f[idx.numpy()] - fit_ref).max())
    print(f"{prefix} P={P}: max|dimg| {img_err:.2e}, rel-L2 {rel:.2e}, max|dfit| {fit_err:.2e}")
    assert img_err <= img_gate and rel <= 5e-3 and fit_err <= fit_gate
    return whole


def test_config3_vqvae_decoder_full_size(engine):
    """BASELINE config 3: `VectorQuantizedVAE_GAN.decoder` (src/inverter/utils_vq_vae/util_model.py:283-289), 16 384
    particles, per-particle fp32 targets, fitness = mean((target - G(z))^2) + eps.
    Measured on B200: max|dimg| 7.8e-3, rel-L2 2.3e-3, max|dfit| 4e-4 (gates: north_star's 2e-2 / 5e-3; fitness 2e-3)."""
    torch.manual_seed(3)
    dec = port.VQDecoderGAN(1, 100, 64).eval()
    port.randomize_bn_stats(dec, torch.Generator().manual_seed(4))
    _check_recon_config(engine, dec, "decoder", 16384, 784, torch.float32, 48, 2e-2, 2e-3)


def test_config4_g256_full_size(engine):
    """BASELINE config 4: G256(z=100, f=64) (extension, oracle port.G256), 8 192 particles, bf16 1x256x256 targets,
    inverter loss with w_ass = 0.  The real layer widths (2048 -> 1024 -> ... -> 64 channels: streamed CTA-pair layers, the
    64 -> 64 epilogue-bound layer, the 128^2 -> 256^2 image layer with the reconstruction epilogue) -- none of which the
    reduced-width test in test_gpu_layers.py exercises."""
    torch.manual_seed(4)
    m = port.G256(100, 1, 64).eval()
    port.initialize_weights(m)
    port.randomize_bn_stats(m, torch.Generator().manual_seed(5))
    whole = _check_recon_config(engine, m, "gen", 8192, 65536, torch.bfloat16, 48, 2e-2, 2e-3)
    # per-layer tcgen05 == SIMT interpreter of the same tables at the real widths, one particle tile
    from psob import _ffi
    P = 128
    z = whole.z[:P].clone()
    outs = {}
    for impl in (_ffi.IMPL_SIMT, _ffi.IMPL_TCGEN05):
        net = engine.GeneratorNet(m.state_dict(), P, impl=impl)
        net.forward(z)
        torch.cuda.synchronize()
        outs[impl] = [net.activation(i).float().cpu() for i in range(len(net.layers) - 1)] + [net.image().cpu()]
        del net
    for i, (s, t) in enumerate(zip(outs[_ffi.IMPL_SIMT], outs[_ffi.IMPL_TCGEN05])):
        # same bf16 inputs per layer only for layer 0; later layers see inputs that differ by re-rounding noise
        err = float((s - t).abs().max()) / max(1.0, float(s.abs().max()))
        assert err <= 1e-2, f"layer {i}: tcgen05 vs SIMT {err:.2e}"


def test_config2_bench_weights_fitness_vs_oracle(engine):
    """The bench's own configuration (initialize_weights: N(0, 0.02) on every conv / BN weight).  MEASURED on B200: the fp32
    oracle's fitness over 48 particles spans 1.8e-7 -- three fp32 ulps of 0.6: with these weights the discriminator's score is
    0.5 for every particle to fp32 resolution, so there is no fitness ORDER to agree on (the reference's own fp32 argmin is
    rounding noise on this config; Spearman against the bf16 engine 0.46) and the global-best index of the bench run carries
    no information.  What can be asserted here is the absolute agreement (measured 6.7e-5, gate 2e-4) and that the engine's
    global best is the strict-< rule applied to ITS fitness vector; ordering / argmin agreement with the oracle is asserted
    where the fitness has a spread to order: torch-default init + randomised BN statistics (test_gpu_swarm.py,
    test_gpu_fused_tail.py, test_config2_default_init_argmin_agrees_with_oracle below)."""
    from src.utils import util_dcgan as M
    torch.manual_seed(42)
    gen, disc = M.Generator(100, 1, 64), M.Discriminator(1, 64)
    M.initialize_weights(gen)
    M.initialize_weights(disc)
    gen, disc = gen.eval(), disc.eval()
    pg, pd_ = port.Generator(100, 1, 64).eval(), port.Discriminator(1, 64).eval()
    pg.load_state_dict(gen.state_dict())
    pd_.load_state_dict(disc.state_dict())
    P = 48
    gnet = engine.GeneratorNet(gen.state_dict(), P)
    dnet = engine.DiscriminatorNet(disc.state_dict(), P)
    sw = engine.SwarmEngine(P, D, generator=gnet, fitness=engine.DScoreFitness(dnet), seed=42)
    sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(42))
    pos0 = sw.pos[:, :D].cpu()
    sw.evaluate_and_best()
    torch.cuda.synchronize()
    fit = sw.fit.cpu().numpy().astype(np.float64)
    ref = port.fitness_dscore(pg, pd_, pos0, D)[0].astype(np.float64)
    err = float(np.abs(fit - ref).max())
    print(f"bench-config fitness: oracle spread {ref.max() - ref.min():.2e}, max|d| {err:.2e}")
    assert err <= 2e-4 and ref.max() - ref.min() <= 1e-5
    f32 = sw.fit.cpu().numpy()
    gval, gidx = port.gbest_scan(f32, np.inf)
    assert int(sw.key) == -1 and float(sw.gbest_val) == gval and torch.equal(sw.gbest_pos[:D].cpu(), pos0[gidx])


def test_config2_default_init_argmin_agrees_with_oracle(engine):
    """Same networks with torch-default init, randomised BN statistics and the discriminator's weights scaled by 3 (as in
    test_gpu_layers.py), so that the fitness has a spread to order (with plain default init it spans 1.9e-3 and the engine
    agrees to 1.2e-4, Spearman 0.989 -- measured).  The engine's bf16 fitness must reproduce the oracle's order: Spearman
    >= 0.98 over 48 particles and the oracle's best particle among the engine's three best."""
    torch.manual_seed(7)
    pg, pd_ = port.Generator(100, 1, 64), port.Discriminator(1, 64)
    port.randomize_bn_stats(pg, torch.Generator().manual_seed(8))
    with torch.no_grad():
        for prm in pd_.parameters():
            if prm.dim() == 4:
                prm.mul_(3.0)
    pg, pd_ = pg.eval(), pd_.eval()
    P = 48
    gnet = engine.GeneratorNet(pg.state_dict(), P)
    dnet = engine.DiscriminatorNet(pd_.state_dict(), P)
    sw = engine.SwarmEngine(P, D, generator=gnet, fitness=engine.DScoreFitness(dnet), seed=42)
    sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(42))
    pos0 = sw.pos[:, :D].cpu()
    sw.evaluate_and_best()
    torch.cuda.synchronize()
    fit = sw.fit.cpu().numpy().astype(np.float64)
    ref = port.fitness_dscore(pg, pd_, pos0, D)[0].astype(np.float64)
    rank = lambda v: np.argsort(np.argsort(v)).astype(np.float64)
    rho = float(np.corrcoef(rank(fit), rank(ref))[0, 1])
    err = float(np.abs(fit - ref).max())
    print(f"default-init fitness: oracle spread {ref.max() - ref.min():.2e}, max|d| {err:.2e}, Spearman {rho:.4f}")
    assert err <= 5e-3 and rho >= 0.98 and ref.max() - ref.min() >= 1e-2
    assert int(np.argmin(ref)) in np.argsort(fit)[:3]
