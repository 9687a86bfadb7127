"""CPU: the pixel-GEMM lowering tables reproduce torch's convolutions (geometry exact in fp32,
bf16 pipeline within the stated tolerance)."""
import os

import numpy as np
import pytest
import torch

from oracle import port
from psob import lowering as L
from psob import nets
from conftest import load_sd


@pytest.fixture
def exact():
    L.ROUND_WEIGHTS = False
    yield
    L.ROUND_WEIGHTS = True


def _z_rows(z, pitch):
    x = np.zeros((z.shape[0], pitch), np.float32)
    x[:, :z.shape[1]] = z
    return x


@pytest.mark.parametrize("m_last", [1, 2, 4])
def test_generator_geometry_exact(golden, exact, m_last):
    sd = load_sd(golden, "gsmall_trained")
    progs = nets.generator_programs(sd, m_last=m_last)
    z = golden["gsmall/z"].reshape(6, 10)
    _, img = nets.emulate_chain(progs, _z_rows(z, progs[0].in_pitch), round_out=False)
    ref = golden["gsmall_trained/out"].reshape(6, 784)
    np.testing.assert_allclose(img, ref, rtol=0, atol=2e-6)
    # exact MAC bookkeeping equals the border-aware count of SURVEY.md §8 (scaled to this net)
    ones = torch.ones(1, 1, 7, 7)
    exact_l2 = int(torch.nn.functional.conv_transpose2d(ones, torch.ones(1, 1, 4, 4), stride=2, padding=1).sum())
    assert progs[1].macs_exact == exact_l2 * 32 * 16


def test_generator_full_size_geometry_and_macs(exact):
    torch.manual_seed(0)
    gen = port.Generator(100, 1, 64).eval()
    port.randomize_bn_stats(gen, torch.Generator().manual_seed(1))
    progs = nets.generator_programs(gen.state_dict())
    z = torch.randn(3, 100, generator=torch.Generator().manual_seed(2))
    with torch.no_grad():
        ref = gen(z.view(3, 100, 1, 1)).reshape(3, 784).numpy()
    _, img = nets.emulate_chain(progs, _z_rows(z.numpy(), progs[0].in_pitch), round_out=False)
    np.testing.assert_allclose(img, ref, rtol=0, atol=5e-6)
    assert [p.macs_exact for p in progs] == [627200, 5537792, 186624]
    assert progs[1].macs_issued == 6422528                                   # nominal implicit-GEMM count
    assert progs[1].mma_n == 256 and progs[0].mma_n == 256 and progs[2].mma_n == (L.BAND_N if L.IMAGE_BAND else 16)


def test_generator_bf16_pipeline_within_tolerance(golden):
    for regime in ("refinit", "default", "trained"):
        sd = load_sd(golden, f"gsmall_{regime}")
        progs = nets.generator_programs(sd)
        z = L.to_bf16_f32(golden["gsmall/z"].reshape(6, 10))
        img_pad, img = nets.emulate_chain(progs, _z_rows(z, progs[0].in_pitch))
        ref = golden[f"gsmall_{regime}/out"].reshape(6, 784)
        assert np.abs(img - ref).max() <= 2e-2                 # north_star bf16 tolerance
        assert np.linalg.norm(img - ref) / np.linalg.norm(ref) <= 5e-3   # SURVEY.md §7 relative gate
        # padded bf16 copy holds the same pixels at (y+1, x+1), zeros elsewhere
        pad = img_pad.reshape(6, L.IMG_PAD_ROWS, L.IMG_PAD_PITCH)
        np.testing.assert_array_equal(pad[:, 1:29, 1:29].reshape(6, 784), L.to_bf16_f32(img))
        assert pad[:, 0].max() == 0 and pad[:, 29:].max() == 0 and pad[:, :, 0].max() == 0 and pad[:, :, 29:].max() == 0


def test_discriminator_and_encoder_geometry_exact(golden, exact):
    x = golden["dsmall/x"]
    xp = np.zeros((6, L.IMG_PAD_ROWS, L.IMG_PAD_PITCH), np.float32)
    xp[:, 1:29, 1:29] = x.reshape(6, 28, 28)
    xp = xp.reshape(6, -1)
    d = nets.discriminator_programs(load_sd(golden, "dsmall"))
    _, score = nets.emulate_chain(d, xp, round_out=False)
    np.testing.assert_allclose(score[:, 0], golden["dsmall/out"].reshape(6), rtol=0, atol=2e-6)
    e = nets.encoder_programs(load_sd(golden, "esmall"))
    _, lat = nets.emulate_chain(e, xp, round_out=False)
    np.testing.assert_allclose(lat, golden["esmall/out"].reshape(6, 10), rtol=0, atol=5e-6)
    # two position-independent banded tiles (+ the bias tile of the constant K-chunk when the bias rides through the MMA)
    assert d[0].b_resident and d[0].wpack.shape == (512 + (256 if d[0].meta.get("ones_cols") else 0), 64)


def test_discriminator_full_size_macs(exact):
    torch.manual_seed(3)
    disc = port.Discriminator(1, 64).eval()
    progs = nets.discriminator_programs(disc.state_dict())
    assert sum(p.macs_exact for p in progs) == 5730688
    x = torch.tanh(torch.randn(2, 1, 28, 28, generator=torch.Generator().manual_seed(4)))
    xp = np.zeros((2, L.IMG_PAD_ROWS, L.IMG_PAD_PITCH), np.float32)
    xp[:, 1:29, 1:29] = x.numpy().reshape(2, 28, 28)
    _, score = nets.emulate_chain(progs, xp.reshape(2, -1), round_out=False)
    with torch.no_grad():
        ref = disc(x).reshape(2).numpy()
    np.testing.assert_allclose(score[:, 0], ref, rtol=0, atol=2e-6)


def test_g256_geometry_small(exact):
    torch.manual_seed(5)
    g = port.G256(z_dim=10, channels_img=1, features_g=4).eval()   # 128..4 channels, 256x256 output
    port.randomize_bn_stats(g, torch.Generator().manual_seed(6))
    progs = nets.g256_programs(g.state_dict())
    z = torch.randn(1, 10, generator=torch.Generator().manual_seed(7))
    with torch.no_grad():
        ref = g(z.view(1, 10, 1, 1)).reshape(1, -1).numpy()
    _, img = nets.emulate_chain(progs, _z_rows(z.numpy(), progs[0].in_pitch), round_out=False)
    np.testing.assert_allclose(img, ref, rtol=0, atol=5e-6)


def test_channel_chunked_jobs(exact):
    """Cout > 64 with m=1 splits into per-chunk jobs with their own bias rows (G256 widths)."""
    rng = np.random.RandomState(0)
    w = rng.randn(64, 128, 4, 4).astype(np.float32) * 0.05
    b = rng.randn(128).astype(np.float32)
    pr = L.lower_convt_k4s2p1("t", w, b, L.ACT_RELU, 3, 3, m=1)
    assert pr.bias.shape[0] == 2 and pr.mma_n == 256
    x = rng.randn(2, 64, 3, 3).astype(np.float32)
    ref = torch.relu(torch.nn.functional.conv_transpose2d(torch.from_numpy(x), torch.from_numpy(w),
                                                          torch.from_numpy(b), stride=2, padding=1))
    xin = np.ascontiguousarray(x.transpose(0, 2, 3, 1)).reshape(2, -1)
    y, _ = L.emulate(pr, xin, round_out=False)
    np.testing.assert_allclose(y.reshape(2, 6, 6, 128).transpose(0, 3, 1, 2), ref.numpy(), rtol=0, atol=1e-5)


def test_program_tables_are_tma_legal():
    torch.manual_seed(0)
    for progs in (nets.generator_programs(port.Generator(100, 1, 64).state_dict()),
                  nets.discriminator_programs(port.Discriminator(1, 64).state_dict())):
        for p in progs:
            assert p.steps[:, 0].min() >= 0 and (p.steps[:, 0] + L.KCHUNK).max() <= p.in_pitch
            assert (p.steps[:, 0] % 8 == 0).all() and p.in_pitch % 8 == 0
            assert (p.steps[:, 1] + p.mma_n).max() <= p.wpack.shape[0]
            assert (p.steps[:, 2] + p.mma_n).max() <= p.nt
            assert p.nt <= 256 and p.mma_n % 16 == 0
            first = p.steps[p.jobs[p.chains[:, 0], 0], 3] & L.STEP_ACCUM
            assert (first == 0).all(), "the first step of every chain must overwrite its accumulator"
            assert p.chains[:, 1].sum() == p.n_jobs and (p.chains[1:, 0] == p.chains[:-1, 0] + p.chains[:-1, 1]).all()
            packed = L.pack_steps(p)
            assert ((packed[:, 1] & 7) >= 1).all() and ((packed[:, 1] & 7) <= 4).all()


def test_head_weights_reproduce_the_score_layer():
    """nets.head_weights (the vector psob_layer_run_dot reduces against) . hidden activation + bias, through the sigmoid,
    equals the score layer's own table emulation (lower_conv_full, src/utils/util_dcgan.py:117-118) on the same input."""
    torch.manual_seed(3)
    disc = port.Discriminator(1, 64)
    progs = nets.discriminator_programs({k: v.detach().numpy() for k, v in disc.state_dict().items()})
    hid, last = progs[1], progs[2]
    assert last.meta["kind"] == "conv_full" and last.meta["out_c"] == 1 and last.in_pitch == hid.out_pitch
    w = nets.head_weights(last)
    assert w.shape == (hid.out_pitch,) and np.array_equal(L.to_bf16_f32(w), w)            # bf16-exact values
    assert np.count_nonzero(w) == 128 * 49                                               # every (pixel, channel) of the 7x7x128 map
    rng = np.random.default_rng(0)
    x = L.to_bf16_f32(rng.standard_normal((5, last.in_pitch)).astype(np.float32))
    _, score = L.emulate(last, x)
    ref = 1.0 / (1.0 + np.exp(-(x.astype(np.float64) @ w.astype(np.float64) + float(last.bias[0, 0]))))
    np.testing.assert_allclose(score[:, 0], ref, atol=1e-6, rtol=0)
    # conv weight layout: w[pix * 128 + ch] = W[0, ch, ky, kx]
    W = disc.state_dict()["disc.3.weight"].numpy()
    np.testing.assert_array_equal(w.reshape(49, 128), L.to_bf16_f32(W[0].reshape(128, 49).T))


def test_vq_grid_decoder_tables_match_reference_geometry(monkeypatch):
    """`VectorQuantizedVAE.decoder` (src/inverter/utils_vq_vae/util_model.py:194-203; two ResBlockBatchNorm on the 7x7 grid,
    in-place first ReLU, identity shortcut as identity ops with cout < channel pitch) lowered to engine tables: the numpy
    emulation of the tables equals the LIVE reference's output (tests/golden/ref_vq.npz) to fp32 noise without bf16
    rounding, and stays within the image gates with it."""
    import os
    from psob import nets
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_vq.npz"))
    sd = {k[len("vqgrid/"):]: g[k] for k in g.files if k.startswith("vqgrid/decoder.")}
    ref = g["vqgrid/out"].reshape(6, -1)
    ops, bufs, dim = nets.vq_grid_decoder_programs(sd)
    img = nets.emulate_vq_grid_decoder(ops, bufs, dim, g["vqgrid/z"])
    assert np.abs(img - ref).max() <= 2e-2 and np.linalg.norm(img - ref) <= 5e-3 * np.linalg.norm(ref)
    monkeypatch.setattr(L, "ROUND_WEIGHTS", False)
    ops, bufs, dim = nets.vq_grid_decoder_programs(sd)
    img = nets.emulate_vq_grid_decoder(ops, bufs, dim, g["vqgrid/z"], round_out=False)
    assert np.abs(img - ref).max() <= 5e-6


@pytest.mark.parametrize("h,wd,cin", [(14, 14, 8), (20, 18, 70), (6, 30, 5)])
def test_tiled_image_band_matches_conv_transpose(monkeypatch, h, wd, cin):
    """The band form for wide images (G256's 128^2 -> 256^2 image layer): tiles of 8 padded rows x 28 columns, cut weight
    tiles at the tile's left / right neighbours, partial tiles at the right border -- exact against torch in fp32."""
    monkeypatch.setattr(L, "ROUND_WEIGHTS", False)
    torch.manual_seed(h * wd)
    m = torch.nn.ConvTranspose2d(cin, 1, 4, 2, 1)
    pr = L.lower_convt_image_band_tiled("t", m.weight.detach().numpy(), m.bias.detach().numpy(), L.ACT_TANH, h, wd)
    x = torch.randn(3, cin, h, wd)
    ref = torch.tanh(m(x)).detach().numpy().reshape(3, -1)
    cp = L.cpitch(cin)
    xin = np.zeros((3, h * wd * cp), np.float32)
    xin.reshape(3, h * wd, cp)[:, :, :cin] = x.permute(0, 2, 3, 1).reshape(3, h * wd, cin).numpy()
    _, y32 = L.emulate(pr, xin, round_out=False)
    assert np.abs(y32 - ref).max() <= 5e-6
    conn = lambda n_in: sum(1 for o in range(2 * n_in) for k in range(4) if (o + 1 - k) % 2 == 0 and 0 <= (o + 1 - k) // 2 < n_in)
    assert pr.macs_exact == conn(h) * conn(wd) * cin and pr.out_pitch == 0 and pr.out32_pitch == 4 * h * wd


def test_pack_tables_reconstructs_every_step_list():
    """The tables the engine uploads store step patterns relative to a per-job base column and share them between jobs
    (`pack_tables`): expanding them again must give back `pack_steps` job by job -- for plain, chained (NEXT ops), banded, tiled-band
    and identity-op layers -- and the wide layers must shrink enough to fit the kernel's shared-memory tables."""
    torch.manual_seed(0)
    g, d = port.Generator(100, 1, 64).eval(), port.Discriminator(1, 64).eval()
    progs = nets.generator_programs({k: v.numpy() for k, v in g.state_dict().items()}) + \
        nets.discriminator_programs({k: v.numpy() for k, v in d.state_dict().items()})
    g256 = port.G256(10, 1, 8).eval()
    progs += nets.g256_programs({k: v.numpy() for k, v in g256.state_dict().items()}, "gen")
    vq = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_vq.npz"))
    ops, _, _ = nets.vq_grid_decoder_programs({k[len("vqgrid/"):]: vq[k] for k in vq.files if k.startswith("vqgrid/decoder.")})
    progs += [op.prog for op in ops]
    for prog in progs:
        full = L.pack_steps(prog)
        steps, jobs = L.pack_tables(prog)
        ref_jobs = L.pack_jobs(prog)
        assert steps.shape[0] <= full.shape[0] and jobs.shape == ref_jobs.shape
        for j in range(jobs.shape[0]):
            sb, sc = int(ref_jobs[j, 0]), int(ref_jobs[j, 1])
            base = (int(jobs[j, 7]) >> 1) * 8
            blk = steps[int(jobs[j, 0]):int(jobs[j, 0]) + sc].copy()
            assert int(jobs[j, 1]) == sc and base % 64 == 0 and (int(jobs[j, 7]) & 1) == int(ref_jobs[j, 7])
            blk[:, 0] += base
            assert np.array_equal(blk, full[sb:sb + sc]), f"{prog.name} job {j}"
            assert np.array_equal(jobs[j, 1:7], ref_jobs[j, 1:7])
    wide = next(p for p in progs if p.meta.get("kind") == "convt_image_band_tiled")
    assert L.pack_tables(wide)[0].shape[0] <= 1024 < wide.n_steps


def test_merged_wide_ops_equal_single_ops():
    """3x3 conv with resident weights (the assessor's stage-1 shape): neighbouring output pixels fed by one input pixel through
    neighbouring taps are merged into ops of 2-3 x mma_n columns (weight tiles of a kernel row stored back to back).  Same
    products, same accumulation order per column: the emulation must agree bit for bit with the single-op tables, the op count
    must drop, and the packed op words must carry the width in bits 28..29."""
    from psob import lowering as L
    rng = np.random.default_rng(0)
    w = (rng.standard_normal((64, 64, 3, 3)) * 0.1).astype(np.float32)
    b = rng.standard_normal(64).astype(np.float32)
    x = L.to_bf16_f32(rng.standard_normal((5, 49 * 64)).astype(np.float32))
    old = L.MERGE_OPS
    try:
        L.MERGE_OPS = True
        pm = L.lower_conv2d("m", w, b, L.ACT_RELU, 7, 7, 1, 1)
        L.MERGE_OPS = False
        pu = L.lower_conv2d("u", w, b, L.ACT_RELU, 7, 7, 1, 1)
    finally:
        L.MERGE_OPS = old
    ym, _ = L.emulate(pm, x)
    yu, _ = L.emulate(pu, x)
    assert np.array_equal(ym, yu)
    n_ops = lambda p: p.n_steps + sum(len(v) for v in p.xops.values())
    assert n_ops(pu) == 361 and n_ops(pm) < 0.65 * n_ops(pu) and pm.n_steps == pu.n_steps
    assert pm.macs_issued == pu.macs_issued                      # a wide op counts its width
    st = L.pack_steps(pm)
    widths = 1 + ((st[:, 2:6].astype(np.int64) >> 28) & 3)
    assert widths.max() == 3 and ((L.pack_steps(pu)[:, 2:6].astype(np.int64) >> 28) & 3).max() == 0
    # a wide op stays inside the job's columns and the weight pack
    for i in range(pm.n_steps):
        for (br, dc, fl) in [tuple(int(v) for v in pm.steps[i][1:4])] + [tuple(int(v) for v in o) for o in pm.xops.get(i, [])]:
            n = pm.mma_n * L.op_width(fl)
            assert dc + n <= pm.nt and br + n <= pm.wpack.shape[0]
