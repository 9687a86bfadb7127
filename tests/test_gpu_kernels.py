"""GPU (-m gpu): libpsob.so kernels vs the oracle (oracle/port.py) and the golden vectors of the
live reference, called through the C ABI (psob._ffi via psob.engine)."""
import math

import numpy as np
import pytest
import torch

from oracle import port

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from psob import engine
    engine.require_cuda()
    return engine


def dev(x, dtype=torch.float32):
    return torch.as_tensor(np.asarray(x), dtype=dtype).cuda().contiguous()


# ---------------------------------------------------------------- update kernel
def test_update_injected_r_matches_reference_move(eng, golden):
    g = golden
    w, wc, ws = (float(v) for v in g["move/w"])
    pos, vel, pb, gb = dev(g["move/pos"]), dev(g["move/vel"]), dev(g["move/pbest"]), dev(g["move/gbest"])
    r1, r2 = dev(g["move/r1"]), dev(g["move/r2"])
    eng.pso_update(pos, vel, pb, gb, r1=r1, r2=r2, w=w, c_cogn=wc, c_soci=ws)
    torch.cuda.synchronize()
    for got, key in ((pos, "move/new_pos"), (vel, "move/new_vel")):
        ref = g[key]
        err = np.abs(got.cpu().numpy().astype(np.float64) - ref)
        assert np.all(err <= 1e-6 * np.maximum(1.0, np.abs(ref)))          # north_star: 1e-6 relative (fp32 state)
    # bit-exact against the same-order fp32 restatement (inputs rounded to fp32 like the device state)
    f32 = lambda k: g[k].astype(np.float32)
    p32, v32 = port.pso_move(f32("move/pos"), f32("move/vel"), f32("move/pbest"), f32("move/gbest"),
                             f32("move/r1"), f32("move/r2"), w, wc, ws, dtype=np.float32)
    assert np.array_equal(pos.cpu().numpy(), p32) and np.array_equal(vel.cpu().numpy(), v32)


@pytest.mark.parametrize("P,D", [(1, 4), (37, 100), (1000, 12), (65536, 100)])
def test_update_philox_bit_exact_and_shard_independent(eng, P, D):
    rng = np.random.RandomState(P)
    pos0 = rng.randn(P, D).astype(np.float32)
    vel0 = rng.randn(P, D).astype(np.float32) * 0.1
    pb0 = (pos0 + 0.3 * rng.randn(P, D)).astype(np.float32)
    gb = rng.randn(D).astype(np.float32)
    seed, it, w, wc, ws = 42, 7, 0.73, 1.496, 1.3
    r1, r2 = port.philox_r1r2(seed, it, 1000, P)
    p_ref, v_ref = port.pso_move(pos0, vel0, pb0, gb, r1, r2, w, wc, ws, dtype=np.float32)
    pos, vel, pb = dev(pos0), dev(vel0), dev(pb0)
    z = torch.zeros(P, 128, dtype=torch.bfloat16, device="cuda")
    eng.pso_update(pos, vel, pb, dev(gb), z=z, w=w, c_cogn=wc, c_soci=ws, seed=seed, iteration=it, index_offset=1000)
    assert np.array_equal(pos.cpu().numpy(), p_ref) and np.array_equal(vel.cpu().numpy(), v_ref)
    assert torch.equal(z[:, :D].float().cpu(), torch.from_numpy(p_ref).bfloat16().float())
    assert float(z[:, D:].abs().max()) == 0.0
    if P >= 1000:       # two shards with global index offsets give the same result as one call
        h = P // 2
        pa, va, pba = dev(pos0[:h]), dev(vel0[:h]), dev(pb0[:h])
        pb_, vb, pbb = dev(pos0[h:]), dev(vel0[h:]), dev(pb0[h:])
        eng.pso_update(pa, va, pba, dev(gb), w=w, c_cogn=wc, c_soci=ws, seed=seed, iteration=it, index_offset=1000)
        eng.pso_update(pb_, vb, pbb, dev(gb), w=w, c_cogn=wc, c_soci=ws, seed=seed, iteration=it, index_offset=1000 + h)
        assert np.array_equal(np.concatenate([pa.cpu().numpy(), pb_.cpu().numpy()]), p_ref)


def test_update_improved_copy_clamps_and_errors(eng):
    from psob import _ffi
    rng = np.random.RandomState(0)
    P, D = 64, 8
    pos0, vel0 = rng.randn(P, D).astype(np.float32) * 3, rng.randn(P, D).astype(np.float32)
    pb0, gb = rng.randn(P, D).astype(np.float32), rng.randn(D).astype(np.float32)
    imp = (rng.random_sample(P) < 0.5)
    r1, r2 = rng.random_sample(P).astype(np.float32), rng.random_sample(P).astype(np.float32)
    pb_eff = np.where(imp[:, None], pos0, pb0)
    p_ref, v_ref = port.pso_move(pos0, vel0, pb_eff, gb, r1, r2, 0.5, 0.8, 0.9, v_max=0.7, lo=-2.0, hi=2.5, dtype=np.float32)
    pos, vel, pb = dev(pos0), dev(vel0), dev(pb0)
    eng.pso_update(pos, vel, pb, dev(gb), improved=dev(imp, torch.uint8), r1=dev(r1), r2=dev(r2), v_max=0.7, lo=-2.0, hi=2.5)
    assert np.array_equal(pb.cpu().numpy(), pb_eff)                       # deferred p_best_pos = pos (util_pso.py:124)
    assert np.array_equal(vel.cpu().numpy(), v_ref) and np.array_equal(pos.cpu().numpy(), p_ref)
    assert np.abs(v_ref).max() <= 0.7 and p_ref.min() >= -2.0 and p_ref.max() <= 2.5
    with pytest.raises(_ffi.PsobError):
        eng.pso_update(dev(np.zeros((4, 6))), dev(np.zeros((4, 6))), dev(np.zeros((4, 6))), dev(np.zeros(6)))   # D % 4
    with pytest.raises(_ffi.PsobError):
        eng.pso_update(pos, vel, pb, dev(gb), r1=dev(r1))                 # r1 without r2


# ---------------------------------------------------------------- personal / global best
def _run_best(eng, fit_seq, D=5, shards=1):
    T, P = fit_seq.shape
    rng = np.random.RandomState(1)
    pos_t = rng.randn(T, P, D).astype(np.float32)
    pbv = torch.full((P,), math.inf, device="cuda")
    pbp = dev(pos_t[0])
    gval, gpos = torch.full((1,), math.inf, device="cuda"), torch.zeros(D, device="cuda")
    hist = torch.zeros(T, 2, device="cuda")
    dyn = torch.zeros(2, dtype=torch.int32, device="cuda")
    out = []
    bounds = [(s * P // shards, (s + 1) * P // shards) for s in range(shards)]
    for t in range(T):
        pos = dev(pos_t[t])
        recs = torch.zeros(shards, 4 + D, device="cuda")
        imp_all = torch.zeros(P, dtype=torch.uint8, device="cuda")
        for s, (a, b) in enumerate(bounds):
            key = torch.zeros(1, dtype=torch.int64, device="cuda")
            fit = dev(fit_seq[t, a:b])
            pv, im = pbv[a:b].clone(), torch.zeros(b - a, dtype=torch.uint8, device="cuda")
            eng.pbest_minloc(fit, pv, im, key, index_offset=a)
            eng.gbest_candidate(key, pos[a:b].contiguous(), pbp[a:b].contiguous(), im, pv, recs[s], index_offset=a)
            pbv[a:b] = pv
            imp_all[a:b] = im
        eng.gbest_apply(recs, shards, D, gval, gpos, hist, dyn, 1.0)
        pbp = torch.where(imp_all.bool()[:, None], pos, pbp)          # what psob_update's deferred copy does
        out.append((pbv.cpu().numpy().copy(), float(gval), gpos.cpu().numpy().copy(), pbp.cpu().numpy().copy()))
    return out, hist.cpu().numpy(), pos_t


@pytest.mark.parametrize("shards", [1, 3])
def test_pbest_gbest_rules_match_reference(eng, golden, shards):
    fit = golden["best/fit"]
    out, hist, pos_t = _run_best(eng, fit, shards=shards)
    for t, (pbv, gval, gpos, pbp) in enumerate(out):
        assert np.array_equal(pbv, golden["best/pbest_val"][t])                    # bit-exact personal bests
        assert gval == golden["best/gbest_val"][t]
        assert hist[t, 0] == golden["best/gbest_val_dummy"][t]
        idx = int(golden["best/gbest_idx"][t])
        assert bool(hist[t, 1]) == (idx >= 0)                                       # equal-to-incumbent never replaces
        if idx >= 0:
            assert np.array_equal(gpos, pbp[idx])                                   # lowest index wins ties
    assert int(hist[:, 1].sum()) == int(golden["best/n_improvements"])


def test_minloc_large_random_and_special_values(eng):
    rng = np.random.RandomState(5)
    P = 300001
    fit = rng.random_sample(P).astype(np.float32) + 0.1
    fit[[17, 200000]] = 0.05
    fit[5] = np.nan
    fit[6] = np.inf
    fit[7] = -1.5                      # generic objectives may go negative (hands_on fun2)
    pbv = torch.full((P,), math.inf, device="cuda")
    imp = torch.zeros(P, dtype=torch.uint8, device="cuda")
    key = torch.zeros(1, dtype=torch.int64, device="cuda")
    eng.pbest_minloc(dev(fit), pbv, imp, key, index_offset=123)
    k = int(key.cpu().numpy().view(np.uint64)[0])
    assert (k & 0xFFFFFFFF) == 123 + 7
    assert imp[5].item() == 0 and imp[6].item() == 0 and math.isinf(pbv[5].item())
    ref_val, ref_idx = port.gbest_scan(np.where(np.isnan(fit), np.inf, fit), np.inf)
    assert ref_idx == 7 and (k >> 32) == int(port.f32_order_key(np.float32(ref_val)))


# ---------------------------------------------------------------- fitness transforms
def test_score_fitness_softmax_rescale_recon(eng, golden):
    rng = np.random.RandomState(2)
    p = rng.random_sample(1000).astype(np.float32)
    for mode in ("optimize_in_training", "optimize_out_training"):
        fit = torch.empty(1000, device="cuda")
        eng.score_fitness(dev(p), fit, mode, thr=0.05, eps=0.1)
        ref = port.score_to_fitness(torch.from_numpy(p), mode, 0.05, 0.1).numpy()
        assert np.array_equal(fit.cpu().numpy(), ref)
    with pytest.raises(ValueError):
        eng.score_fitness(dev(p), torch.empty(1000, device="cuda"), "bogus")
    logits = rng.randn(257, 8).astype(np.float32) * 3
    prob = torch.empty(257, device="cuda")
    eng.softmax_select(dev(logits), 3, prob)
    ref = torch.softmax(torch.from_numpy(logits), dim=1)[:, 3].numpy()
    np.testing.assert_allclose(prob.cpu().numpy(), ref, rtol=2e-6, atol=1e-8)
    img = golden["fit_multi/img"].reshape(-1, 784)
    out = torch.empty(img.shape, device="cuda")
    eng.rescale_minmax(dev(img), out)
    ref = port.rescale_minmax_per_image(torch.from_numpy(img)).numpy()
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=0, atol=1e-6)
    np.testing.assert_allclose(out.cpu().numpy()[-1], golden["fit_multi/img_rescaled"].reshape(-1, 784)[-1], rtol=0, atol=1e-6)
    for n in (784, 65536):
        x = np.tanh(rng.randn(33, n)).astype(np.float32)
        t = np.tanh(rng.randn(33, n)).astype(np.float32)
        ass = rng.random_sample(33).astype(np.float32)
        fit = torch.empty(33, device="cuda")
        eng.recon_fitness(dev(x), dev(t), fit, dev(ass), w_ass=0.7, w_rec=1.3, eps=0.1)
        ref = 0.7 * ass + 1.3 * ((t.astype(np.float64) - x) ** 2).mean(axis=1) + 0.1
        np.testing.assert_allclose(fit.cpu().numpy(), ref, rtol=2e-6)
        eng.recon_fitness(dev(x), dev(t).bfloat16(), fit, None, w_rec=1.0, eps=0.1)
        tb = torch.from_numpy(t).bfloat16().float().numpy().astype(np.float64)
        np.testing.assert_allclose(fit.cpu().numpy(), ((tb - x) ** 2).mean(axis=1) + 0.1, rtol=2e-6)


def test_pack_image_layout(eng):
    rng = np.random.RandomState(4)
    img = np.tanh(rng.randn(5, 784)).astype(np.float32)
    out = torch.full((5, 31 * 32), 7.0, dtype=torch.bfloat16, device="cuda")
    eng.pack_image(dev(img), out)
    pad = out.float().cpu().numpy().reshape(5, 31, 32)
    assert np.array_equal(pad[:, 1:29, 1:29].reshape(5, 784), torch.from_numpy(img).bfloat16().float().numpy())
    pad[:, 1:29, 1:29] = 0
    assert np.abs(pad).max() == 0
