"""GPU (-m gpu): the fused fitness tail (psob_fitness_tail + psob_gbest_local / psob_gbest_record) against the unfused
chain of separate kernels and against the oracle's rules; the reconstruction loss accumulated in the generator's last-layer
epilogue (psob_layer_run_recon) against the materialised-image form; regressions for the round-1 advisor findings."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

from oracle import port

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from psob import engine
    engine.require_cuda()
    return engine


def _models(seed=0):
    torch.manual_seed(seed)
    gen = port.Generator(100, 1, 64)
    disc = port.Discriminator(1, 64)
    port.randomize_bn_stats(gen, torch.Generator().manual_seed(seed + 1))
    return gen.eval(), disc.eval()


def _dscore_swarm(engine, gen, disc, P, seed=42, fused=True, **kw):
    gnet = engine.GeneratorNet(gen.state_dict(), P)
    dnet = engine.DiscriminatorNet(disc.state_dict(), P)
    sw = engine.SwarmEngine(P, 100, generator=gnet, fitness=engine.DScoreFitness(dnet), w_inertia=0.73, w_cogn=1.496,
                            w_soci=1.496, seed=seed, **kw)
    sw.fused_tail = fused
    return sw


@pytest.mark.parametrize("P", [300, 4096])
def test_fused_step_equals_unfused_chain_bit_for_bit(engine, P):
    """dot_finish -> score_fitness -> memset + pbest_minloc -> gbest_candidate -> gbest_apply (6 graph nodes) versus
    fitness_tail -> gbest_local (2 launches): same arithmetic in the same order, so every state tensor is identical."""
    gen, disc = _models(1)
    a = _dscore_swarm(engine, gen, disc, P, fused=False)
    b = _dscore_swarm(engine, gen, disc, P, fused=True)
    for sw in (a, b):
        sw.inertia_decay = 0.99
        sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(5))
    for _ in range(5):
        a.step()
        b.step()
    torch.cuda.synchronize()
    for name in ("fit", "pos", "vel", "pbest_pos", "pbest_val", "gbest_pos", "gbest_val", "improved", "hist"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    assert a.iteration_state() == b.iteration_state() and int(b.key) == -1       # key re-armed for the next iteration


def test_fitness_tail_sources_and_rules_vs_oracle(engine):
    """Every score source of psob_fitness_tail against the oracle's restatement of Discovery.fitness /
    fitness_pso_inverter (src/pso/util_discovery.py:70-80,97-101) and the strict-< rules (src/pso/util_pso.py:122-151),
    including ties, NaN, +-0.0 and an equal-to-incumbent candidate."""
    from psob import _ffi
    P = 1000
    g = torch.Generator().manual_seed(3)
    dev = "cuda"
    pbest0 = torch.full((P,), math.inf)
    pbest0[::3] = torch.rand(P, generator=g)[::3] * 0.5 + 0.1

    def run(fields, pbest):
        fit = torch.zeros(P, device=dev)
        pb = pbest.clone().to(dev)
        imp = torch.zeros(P, dtype=torch.uint8, device=dev)
        key = torch.full((1,), -1, dtype=torch.int64, device=dev)
        a = _ffi.FitnessTailArgs(fit=fit.data_ptr(), pbest_val=pb.data_ptr(), improved=imp.data_ptr(), P=P, index_offset=7,
                                 key=key.data_ptr(), **fields)
        _ffi.check(_ffi.lib().psob_fitness_tail(C.byref(a), torch.cuda.current_stream().cuda_stream), "tail")
        torch.cuda.synchronize()
        return fit.cpu().numpy(), pb.cpu().numpy(), imp.cpu().numpy(), int(key.cpu().numpy().view(np.uint64)[0])

    def check_rules(fit, pb, imp, key, pbest):
        pb_ref, imp_ref = port.pbest_update(fit, pbest.numpy())
        assert np.array_equal(pb, pb_ref, equal_nan=True) and np.array_equal(imp.astype(bool), imp_ref)
        keys = port.pack_minloc_key(pb_ref, np.arange(P) + 7)
        assert key == int(keys.min())

    # (a) ready probability, both modes
    score = torch.rand(P, generator=g)
    for mode, mname in ((_ffi.MODE_OUT_TRAINING, "optimize_out_training"), (_ffi.MODE_IN_TRAINING, "optimize_in_training")):
        s_dev = score.to(dev)
        fit, pb, imp, key = run(dict(score=s_dev.data_ptr(), score_stride=1, mode=mode, thr=0.05, eps=0.1), pbest0)
        ref = port.score_to_fitness(score, mname, thr=0.05, eps=0.1).numpy()
        assert np.array_equal(fit, ref)
        check_rules(fit, pb, imp, key, pbest0)
    # (b) logits -> softmax pick
    logits = torch.randn(P, 8, generator=g) * 3
    l_dev = logits.to(dev)
    fit, pb, imp, key = run(dict(logits=l_dev.data_ptr(), n_class=8, class_idx=5, mode=_ffi.MODE_OUT_TRAINING, thr=0.0, eps=0.1), pbest0)
    ref = port.score_to_fitness(torch.softmax(logits, dim=1, dtype=torch.float32)[:, 5], "optimize_out_training").numpy()
    np.testing.assert_allclose(fit, ref, atol=2e-6, rtol=0)
    check_rules(fit, pb, imp, key, pbest0)
    # (c) fused-head partial sums + bias + sigmoid == psob_dot_finish + psob_score_fitness, bit for bit
    part = torch.randn(5, P, generator=g).to(dev)
    fit, pb, imp, key = run(dict(part=part.data_ptr(), n_slots=5, bias=0.25, act=4, mode=_ffi.MODE_OUT_TRAINING, thr=0.0, eps=0.1), pbest0)
    sc = torch.empty(P, device=dev)
    f2 = torch.empty(P, device=dev)
    _ffi.check(_ffi.lib().psob_dot_finish(part.data_ptr(), 5, P, 0.25, 4, sc.data_ptr(), torch.cuda.current_stream().cuda_stream), "finish")
    engine.score_fitness(sc, f2, "optimize_out_training")
    assert np.array_equal(fit, f2.cpu().numpy())
    check_rules(fit, pb, imp, key, pbest0)
    # (d) ready fitness with special values: ties on the minimum, NaN, -0.0 / +0.0, negative values, equal to the incumbent
    f = torch.randn(P, generator=g)
    f[10] = f[500] = f.min() - 1.0                      # tie on the global minimum: the lower index must win
    f[3] = float("nan")
    f[20], f[21] = -0.0, 0.0
    pb_special = pbest0.clone()
    pb_special[30] = f[30]                               # equal never improves
    f_dev = f.to(dev)
    fit, pb, imp, key = run(dict(fit_in=f_dev.data_ptr()), pb_special)
    check_rules(f.numpy(), pb, imp, key, pb_special)
    assert (key & 0xFFFFFFFF) == 10 + 7 and imp[3] == 0 and imp[30] == 0
    # (e) reconstruction partial sums on top of a score: w_ass * fitness + w_rec * sum / n_pix + eps (eps twice, :101)
    rpart = torch.rand(7, P, generator=g).to(dev)
    s_dev = score.to(dev)
    fit, pb, imp, key = run(dict(score=s_dev.data_ptr(), score_stride=1, mode=_ffi.MODE_OUT_TRAINING, thr=0.0, eps=0.1,
                                 recon_part=rpart.data_ptr(), n_recon_slots=7, n_pix=784, w_ass=0.7, w_rec=1.3, eps_rec=0.1), pbest0)
    mse = rpart.cpu().numpy().astype(np.float64).sum(0) / 784
    ref = 0.7 * port.score_to_fitness(score, "optimize_out_training").numpy().astype(np.float64) + 1.3 * mse + 0.1
    np.testing.assert_allclose(fit, ref, rtol=2e-6, atol=0)
    check_rules(fit, pb, imp, key, pbest0)
    # argument errors
    bad = _ffi.FitnessTailArgs(fit=None, pbest_val=1, improved=1, P=P, key=1, score=1, score_stride=1, fit_in=1)
    assert _ffi.lib().psob_fitness_tail(C.byref(bad), None) < 0


def test_recon_epilogue_equals_materialised_image_form(engine):
    """Config 3 shape (VQ-VAE-GAN decoder, fp32 and bf16 targets): the squared error accumulated inside the image band's
    epilogue == psob_recon_fitness on the stored fp32 image, up to the summation order."""
    P = 700
    torch.manual_seed(5)
    dec = port.VQDecoderGAN(1, 100, 64).eval()
    for tdt in (torch.float32, torch.bfloat16):
        tgt = (torch.rand(P, 784, generator=torch.Generator().manual_seed(6)) * 2 - 1).cuda().to(tdt)
        sws = []
        for fused in (False, True):
            gnet = engine.GeneratorNet(dec.state_dict(), P, prefix="decoder")
            sw = engine.SwarmEngine(P, 100, generator=gnet, fitness=engine.ReconFitness(tgt, w_rec=1.0, eps=0.1), seed=1,
                                    need_image_f32=not fused)
            sw.fused_tail = fused
            sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(3))
            sw.step()
            sws.append(sw)
        torch.cuda.synchronize()
        a, b = sws
        assert not b.write_f32                          # the fused form never writes the fp32 image
        np.testing.assert_allclose(b.fit.cpu().numpy(), a.fit.cpu().numpy(), rtol=3e-6, atol=0)
        assert torch.equal(a.gen.image_padded_bf16(), b.gen.image_padded_bf16())
    # with an assessor (D-score) in front: w_ass * fitness + w_rec * mse + eps
    gen, disc = _models(2)
    tgt = (torch.rand(P, 784, generator=torch.Generator().manual_seed(8)) * 2 - 1).cuda()
    outs = []
    for fused in (False, True):
        gnet = engine.GeneratorNet(gen.state_dict(), P)
        dnet = engine.DiscriminatorNet(disc.state_dict(), P)
        fit = engine.ReconFitness(tgt, w_rec=1.3, eps=0.1, assessor=engine.DScoreFitness(dnet), w_ass=0.7)
        sw = engine.SwarmEngine(P, 100, generator=gnet, fitness=fit, seed=1, need_image_f32=not fused)
        sw.fused_tail = fused
        sw.init_particles(generator=torch.Generator(device="cuda").manual_seed(3))
        pos0 = sw.pos[:, :100].cpu()
        sw.step()
        outs.append(sw.fit.cpu().numpy())
    np.testing.assert_allclose(outs[1], outs[0], rtol=3e-6, atol=0)
    ass = port.fitness_dscore(gen, disc, pos0[:32], 100)[0]
    ref, _ = port.fitness_inverter(gen, pos0[:32], tgt[:32].cpu().view(32, 1, 28, 28), 100, w_ass=0.7, w_rec=1.3, assessor_fitness=ass)
    np.testing.assert_allclose(outs[1][:32], ref, atol=5e-3, rtol=0)       # measured 1.1e-3 (bf16 networks vs fp32 oracle)


def test_shared_generator_net_keeps_its_fp32_image(engine):
    """Round-1 advisor finding: a swarm with a D-score assessor skips the fp32 image store; that must be a per-call
    choice, so Generator.forward / particles_to_img_batch on the SAME cached net (batch == P) still return fresh images."""
    from src.pso import util_pso, util_discovery
    from src.utils import util_dcgan
    torch.manual_seed(0)
    gen = util_dcgan.Generator(100, 1, 64).cuda().eval()
    dsc = util_dcgan.Discriminator(1, 64).cuda().eval()
    P = 256
    disc = util_discovery.Discovery(iid_class=None, model_gan=gen, model_cnn=dsc, device="cuda:0")
    sw = util_pso.Swarm("/tmp/psob_unused", disc, num_particles=P, n_iterations=3, dim_space=100, device="cuda:0",
                        compute_diversity=False, record_history=False)
    sw.optimize(early_stopping=False)
    assert not sw.engine.write_f32 and sw.engine.gen is gen.engine_net(P)
    z = torch.randn(P, 100, 1, 1, device="cuda", generator=torch.Generator(device="cuda").manual_seed(9))
    img = gen(z)
    pg = port.Generator(100, 1, 64).eval()
    pg.load_state_dict({k: v.cpu() for k, v in gen.state_dict().items()})
    with torch.no_grad():
        ref = pg(z.cpu())
    assert float((img.cpu() - ref).abs().max()) <= 2e-2
    assert float(img.abs().max()) > 0                                       # not the zero-initialised workspace
    resc, img2 = disc.particles_to_img_batch(z.view(P, 100))
    assert torch.equal(img2, img) and float(resc.max()) == 1.0


def test_swarm_coefficient_setters_reach_the_engine(engine):
    """Round-1 advisor findings: update_inertia() / `swarm.w_inertia = ...` must change the weight the move kernel reads;
    calling the personal best twice before a move must not lose an improved position; a second optimize() starts clean;
    running past the recorder's capacity raises instead of overwriting the last row."""
    from src.pso import util_pso, util_discovery
    from src.utils import util_dcgan
    torch.manual_seed(0)
    gen = util_dcgan.Generator(100, 1, 64).cuda().eval()
    dsc = util_dcgan.Discriminator(1, 64).cuda().eval()
    disc = util_discovery.Discovery(iid_class=None, model_gan=gen, model_cnn=dsc, device="cuda:0")
    P = 128
    sw = util_pso.Swarm("/tmp/psob_unused", disc, num_particles=P, n_iterations=4, dim_space=100, device="cuda:0",
                        w_inertia=0.5, compute_diversity=False)
    sw.init_particles()
    # the reference-style hand-written loop
    sw.update_personal_best(iteration=1)
    sw.update_global_best()
    sw.update_inertia()
    assert abs(sw.engine.iteration_state()[1] - 0.495) < 1e-7
    pos0, vel0 = sw.engine.pos.cpu().numpy().copy(), sw.engine.vel.cpu().numpy().copy()
    pb0 = pos0.copy()                                    # first iteration: every particle improved
    gb = sw.engine.gbest_pos.cpu().numpy().copy()
    sw.w_cogn, sw.w_soci = 1.1, 1.2
    sw.move_particles()
    r1, r2 = port.philox_r1r2(42, 1, 0, P)
    p_ref, v_ref = port.pso_move(pos0, vel0, pb0, gb, r1, r2, np.float32(0.495), 1.1, 1.2, dtype=np.float32)
    assert np.array_equal(sw.engine.pos.cpu().numpy(), p_ref) and np.array_equal(sw.engine.vel.cpu().numpy(), v_ref)
    # personal best twice without a move: the improved positions of the first call survive
    sw.update_personal_best(iteration=2)
    sw.update_global_best()
    improved = sw.engine.improved.cpu().numpy().astype(bool)
    pos_now = sw.engine.pos.cpu().numpy()
    sw.update_personal_best(iteration=2)
    sw.update_global_best()
    pb = sw.engine.pbest_pos.cpu().numpy()
    assert improved.any() and np.array_equal(pb[improved], pos_now[improved])
    assert np.array_equal(sw.swarm[int(np.flatnonzero(improved)[0])].p_best_pos, pos_now[improved][0, :100].astype(np.float64))
    # a second optimize() on the same object restarts the host mirrors
    sw.optimize(early_stopping=False)
    n1 = len(sw.g_best_val_dummy)
    sw.optimize(early_stopping=False)
    assert n1 == 4 and len(sw.g_best_val_dummy) == 4 and sw._hist_len == 5
    # the recorder is never overrun silently
    from psob import _ffi
    with pytest.raises(_ffi.PsobError):
        sw.engine.run(2, use_graph=False)
