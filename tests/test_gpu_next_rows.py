"""GPU (-m gpu): the "next" rows of the hot-path scope through the C-ABI --
`vq` nearest-codebook search (src/inverter/utils_vq_vae/util_function.py:4-23) and the swarm
diversity `Swarm.mse` (src/pso/util_pso.py:76-86) -- against the oracle and the golden vectors."""
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


def _check_indices(idx, x, cb, ref_idx):
    """Index work is bit-exact wherever the minimum is unique in fp32: a differing index is accepted only when
    the two candidates' exact (fp64) distances agree to 1e-5 relative (the addmm form cancels ~6 digits)."""
    bad = np.nonzero(idx != ref_idx)[0]
    x64, c64 = x.astype(np.float64), cb.astype(np.float64)
    for p in bad:
        d_a = ((x64[p] - c64[idx[p]]) ** 2).sum()
        d_b = ((x64[p] - c64[ref_idx[p]]) ** 2).sum()
        assert abs(d_a - d_b) <= 1e-5 * max(d_a, d_b, (x64[p] ** 2).sum()), (p, idx[p], ref_idx[p], d_a, d_b)
    return len(bad)


def test_vq_matches_golden_indices(engine, golden):
    x, cb = golden["vq/x"], golden["vq/codebook"]
    idx = engine.vq_nearest(torch.from_numpy(x).cuda(), torch.from_numpy(cb).cuda())
    assert idx.dtype == torch.int64
    assert np.array_equal(idx.cpu().numpy(), golden["vq/indices"])


@pytest.mark.parametrize("P,D,K", [(1, 1, 1), (7, 3, 2), (1000, 10, 512), (4096, 256, 512), (333, 64, 1500), (50, 700, 100)])
def test_vq_vs_oracle_shapes(engine, P, D, K):
    g = torch.Generator().manual_seed(P * 31 + D)
    x, cb = torch.randn(P, D, generator=g), torch.randn(K, D, generator=g)
    ref = port.vq_nearest(x, cb).numpy()
    idx, dist = engine.vq_nearest(x.cuda(), cb.cuda(), return_distance=True)
    n_bad = _check_indices(idx.cpu().numpy(), x.numpy(), cb.numpy(), ref)
    assert n_bad <= max(1, P // 200)
    exact = ((x.double()[:, None, :] - cb.double()[None, :, :]) ** 2).sum(-1).min(1).values if P * K * D < 5e7 else None
    if exact is not None:
        scale = (x.double() ** 2).sum(1) + (cb.double() ** 2).sum(1).max()
        assert ((dist.cpu().double() - exact).abs() <= 2e-6 * scale).all()


def test_vq_ties_take_lowest_index_and_leading_shape(engine):
    cb = torch.randn(40, 16, generator=torch.Generator().manual_seed(3))
    cb[33] = cb[5]
    cb[20] = cb[5]
    x = cb[[5, 20, 33, 7]].clone().view(2, 2, 16)
    idx = engine.vq_nearest(x.cuda(), cb.cuda())
    assert idx.shape == (2, 2) and idx.cpu().tolist() == [[5, 5], [5, 7]]
    assert engine.vq_nearest(torch.empty(0, 16).cuda(), cb.cuda()).shape == (0,)


def test_vq_mirror_module_has_reference_signature(engine, golden):
    from src.inverter.utils_vq_vae.util_function import vq
    x = torch.from_numpy(golden["vq/x"]).cuda().requires_grad_(True)
    idx = vq(x, torch.from_numpy(golden["vq/codebook"]).cuda())
    assert not idx.requires_grad and np.array_equal(idx.cpu().numpy(), golden["vq/indices"])


@pytest.mark.parametrize("P,D", [(2, 1), (5, 100), (64, 100), (65, 7), (1000, 100), (2048, 2)])
def test_mean_pairwise_distance_vs_oracle(engine, P, D):
    """fp32 differences and square roots, fp64 sum: 2e-6 relative to the fp64 oracle."""
    pos = torch.randn(P, D, generator=torch.Generator().manual_seed(P + D))
    ref = port.mean_pairwise_distance(pos.numpy())
    got = engine.mean_pairwise_distance(pos.cuda())
    assert abs(got - ref) <= 2e-6 * ref
    # a strided view (the engine's padded position rows) gives the same number (the order of the per-block
    # atomicAdd(double) is not fixed, so to 1e-12 rather than bit for bit)
    wide = torch.zeros(P, D + 3)
    wide[:, :D] = pos
    assert abs(engine.mean_pairwise_distance(wide.cuda()[:, :D]) - got) <= 1e-12 * got


def test_mean_pairwise_distance_degenerate(engine):
    assert np.isnan(engine.mean_pairwise_distance(torch.zeros(1, 4).cuda()))
    assert engine.mean_pairwise_distance(torch.ones(300, 4).cuda()) == 0.0
    # golden: the reference loop's final swarm (P = 8) and its own mean_mse
    with pytest.raises(Exception):
        engine.mean_pairwise_distance(torch.zeros(4, 4, dtype=torch.float64).cuda())


def test_mean_pairwise_distance_matches_reference_loop_value(engine, golden):
    pos = golden["loop/final_pos"]
    got = engine.mean_pairwise_distance(torch.from_numpy(pos).float().cuda())
    assert abs(got - golden["loop/mean_mse"][-1]) <= 2e-6 * got
