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


@pytest.fixture(scope="module")
def golden_vq():
    import os
    path = os.path.join(os.path.dirname(__file__), "golden", "ref_vq.npz")
    with np.load(path) as z:
        return {k: z[k] for k in z.files}


def test_vq_grid_decoder_pso_weights_and_snapped_swarm(golden_vq):
    """SURVEY section 8 rows a20 / f4 against the live reference's outputs (oracle/gen_golden_vq.py):
    `VectorQuantizedVAE.decoder` (two ResBlockBatchNorm on the 7x7 grid, in-place first ReLU) and `decode(latents)` through a
    codebook set by `pso_weights`; tcgen05 == numpy emulation of the tables; the VQ-snapped variant of config 3 --
    `decode_snapped` and a SwarmEngine with `snap_to_codebook` -- reproduces `vq`'s indices and decodes the selected rows."""
    import pandas as pd
    from psob import engine, nets, _ffi
    from src.inverter.utils_vq_vae import util_model as U
    g = golden_vq
    m = U.VectorQuantizedVAE(1, 100, 256, pd.DataFrame(g["vqgrid/codebook"]))
    assert np.array_equal(m.codebook.embedding.weight.detach().numpy(), g["vqgrid/codebook"])           # pso_weights
    sd = {k[len("vqgrid/"):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("vqgrid/decoder.")}
    m.load_state_dict(sd, strict=False)
    m = m.cuda().eval()
    out = m.decoder(torch.from_numpy(g["vqgrid/z"]).cuda()).cpu().numpy()
    ref = g["vqgrid/out"]
    err, rel = float(np.abs(out - ref).max()), float(np.linalg.norm(out - ref) / np.linalg.norm(ref))
    print(f"VectorQuantizedVAE.decoder: max|d| {err:.2e}, rel-L2 {rel:.2e}")
    assert err <= 2e-2 and rel <= 5e-3                        # measured 8.3e-3 / 4.0e-3 (six bf16 layers)
    out = m.decode(torch.from_numpy(g["vqgrid/latents"]).cuda()).cpu().numpy()
    ref = g["vqgrid/decode_out"]
    assert np.abs(out - ref).max() <= 2e-2 and np.linalg.norm(out - ref) <= 5e-3 * np.linalg.norm(ref)
    # engine == numpy emulation of the same tables (re-rounding noise only)
    ops, bufs, dim = nets.vq_grid_decoder_programs({k: v.numpy() for k, v in sd.items()})
    emu = nets.emulate_vq_grid_decoder(ops, bufs, dim, g["vqgrid/z"])
    got = m.decoder(torch.from_numpy(g["vqgrid/z"]).cuda()).cpu().numpy().reshape(6, -1)
    assert np.abs(got - emu).max() <= 4e-3
    with pytest.raises(_ffi.PsobError):
        m(torch.zeros(1, 1, 28, 28).cuda())
    # VQ-snapped config 3
    m2 = U.VectorQuantizedVAE_GAN(1, 100, 64, pd.DataFrame(g["vqsnap/codebook"]), features_g=16, features_d=16)
    m2.load_state_dict({k[len("vqsnap/"):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("vqsnap/decoder.")}, strict=False)
    m2 = m2.cuda().eval()
    x = torch.from_numpy(g["vqsnap/x"]).cuda()
    idx, img = m2.decode_snapped(x)
    assert np.array_equal(idx.cpu().numpy(), g["vqsnap/idx"])
    ref = g["vqsnap/out"]
    assert np.abs(img.cpu().numpy() - ref).max() <= 2e-2 and np.linalg.norm(img.cpu().numpy() - ref) <= 5e-3 * np.linalg.norm(ref)
    P = 12
    gsd = {"decoder." + k: v for k, v in m2.decoder.state_dict().items()}
    gnet = engine.GeneratorNet(gsd, P, prefix="decoder")
    tgt = torch.zeros(P, 784, device="cuda")
    sw = engine.SwarmEngine(P, 100, generator=gnet, fitness=engine.ReconFitness(tgt), seed=1, snap_to_codebook=m2.codebook.embedding.weight.detach())
    sw.init_particles(pos=x, vel=torch.zeros_like(x))
    assert np.array_equal(sw.code_idx.cpu().numpy(), g["vqsnap/idx"])
    z_ref = torch.from_numpy(g["vqsnap/codebook"][g["vqsnap/idx"]]).to(torch.bfloat16)
    assert torch.equal(sw.z[:, :100].cpu(), z_ref)
    sw.step()
    torch.cuda.synchronize()
    # fitness was evaluated on the SNAPPED codes (targets are 0: fit = mean(img^2) + eps); the state moved continuously
    fit_ref = (ref.reshape(P, -1) ** 2).mean(1) + 0.1
    np.testing.assert_allclose(sw.fit.cpu().numpy(), fit_ref, atol=5e-3, rtol=0)
    from oracle import port
    idx2 = port.vq_nearest(sw.pos[:, :100].cpu(), torch.from_numpy(g["vqsnap/codebook"]))
    assert np.array_equal(sw.code_idx.cpu().numpy(), idx2.numpy())
