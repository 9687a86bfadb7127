"""GPU (-m gpu): the reference's ResNet assessor (src/pso/util_cnn.py:81-190) on the pixel-GEMM engine, through the
C-ABI, against the reference's golden logits, the engine's numpy emulation, and `Discovery.fitness`."""
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


def _model(g, name):
    *layers, n_class, seed = (int(v) for v in g[f"{name}/meta"])
    m = port.Resnet(layers, 1, list(range(n_class)) if n_class > 2 else [0, 1], n_class).eval()
    port.seeded_fill(m, seed)
    return m


@pytest.mark.parametrize("name", ["r1111_c8", "r2121_c3", "r50_c2", "r50_c8"])
def test_assessor_logits_vs_reference_golden(engine, golden_resnet, name):
    """bf16 operands / fp32 accumulation over up to 53 layers vs the fp32 reference: |dlogit| <= 3 % of the logit
    range, softmax probabilities within 3e-2."""
    g = golden_resnet
    m = _model(g, name)
    x = torch.from_numpy(g["x"]).reshape(-1, 784).cuda()
    net = engine.AssessorNet(m.state_dict(), x.shape[0])
    got = net.forward(x).cpu().numpy()
    ref = g[f"{name}/logits"]
    assert np.abs(got - ref).max() <= 0.03 * np.abs(ref).max()
    sm = lambda v: torch.softmax(torch.from_numpy(v), 1).numpy()
    assert np.abs(sm(got) - sm(ref)).max() <= 3e-2


def test_assessor_matches_table_emulation_and_simt(engine, golden_resnet):
    """The tcgen05 path, the SIMT interpreter and the numpy emulation execute the same tables: logits agree to
    fp32 summation-order noise amplified by bf16 re-rounding of every activation over the depth of the
    network (measured 0.5-1 % of the logit range; gate 2 %), P = 300 crosses particle tiles."""
    from psob import nets, _ffi
    g = golden_resnet
    m = _model(g, "r2121_c3")
    gen = torch.Generator().manual_seed(3)
    x = torch.rand(300, 784, generator=gen)
    x[7] = 0.0                                   # a constant image
    ops, bufs = nets.assessor_programs(m.state_dict())
    emu = nets.emulate_assessor(ops, bufs, x.numpy())
    tc = engine.AssessorNet(m.state_dict(), 300).forward(x.cuda()).cpu().numpy()
    # the skip connection as MMA steps (default) and as an epilogue residual operand compute the same sums
    epi = engine.AssessorNet(m.state_dict(), 300, shortcut="epilogue").forward(x.cuda()).cpu().numpy()
    assert np.abs(epi - tc).max() <= 2e-2 * np.abs(tc).max()
    simt = engine.AssessorNet(m.state_dict(), 300, impl=_ffi.IMPL_SIMT).forward(x.cuda()).cpu().numpy()
    scale = np.abs(emu).max()
    assert np.abs(tc - emu).max() <= 2e-2 * scale and np.abs(simt - emu).max() <= 2e-2 * scale
    with torch.no_grad():
        ref = m(x.view(-1, 1, 28, 28)).numpy()
    assert np.abs(tc - ref).max() <= 0.03 * np.abs(ref).max()


def test_stem_pack_rescale_and_maxpool_bit_exact(engine):
    from psob import lowering as L, _ffi
    gen = torch.Generator().manual_seed(9)
    img = torch.randn(37, 784, generator=gen)
    m = port.Resnet([1, 1, 1, 1], 1, [0, 1], 2).eval()
    net = engine.AssessorNet(m.state_dict(), 37)
    net.pack(img.cuda(), rescale=True, keep_rescaled=True)
    ref = port.rescale_minmax_per_image(img.view(37, 1, 28, 28)).view(37, 784)
    assert torch.equal(net.rescaled.cpu(), ref)                                    # same fp32 operations, bit for bit
    packed = net.buf["img"].float().cpu().numpy()
    assert np.array_equal(packed, L.pad_image_stem(L.to_bf16_f32(ref.numpy())))
    # max pool: bf16 in, bf16 out, exact
    x = torch.randn(37, 14 * 14 * 64, generator=gen).bfloat16()
    y = torch.zeros(37, 7 * 7 * 64, dtype=torch.bfloat16, device="cuda")
    xc = x.cuda()
    _ffi.check(_ffi.lib().psob_maxpool_nhwc(xc.data_ptr(), xc.stride(0), 37, 14, 14, 64, 3, 2, 1, y.data_ptr(), y.stride(0), 0, None), "maxpool")
    ref = torch.nn.functional.max_pool2d(x.float().view(37, 14, 14, 64).permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1).reshape(37, -1)
    assert torch.equal(y.float().cpu(), ref)
    # tile-major output (the engine's internal layout) holds the same values
    yt = torch.zeros(128, 7 * 7 * 64, dtype=torch.bfloat16, device="cuda")
    _ffi.check(_ffi.lib().psob_maxpool_nhwc(xc.data_ptr(), xc.stride(0), 37, 14, 14, 64, 3, 2, 1, yt.data_ptr(), yt.stride(0), 1, None), "maxpool")
    assert torch.equal(engine.untile(yt, 37).float().cpu(), ref)
    # the general kernel (any window): 2x2 stride 1 pad 1 on a 7x5 map with 128 channels, into a wider output row
    x = torch.randn(37, 7 * 5 * 128, generator=gen).bfloat16()
    xc = x.cuda()
    y = torch.zeros(37, 8 * 6 * 128 + 64, dtype=torch.bfloat16, device="cuda")
    _ffi.check(_ffi.lib().psob_maxpool_nhwc(xc.data_ptr(), xc.stride(0), 37, 7, 5, 128, 2, 1, 1, y.data_ptr(), y.stride(0), 0, None), "maxpool")
    ref = torch.nn.functional.max_pool2d(x.float().view(37, 7, 5, 128).permute(0, 3, 1, 2), 2, 1, 1).permute(0, 2, 3, 1).reshape(37, -1)
    assert torch.equal(y[:, :8 * 6 * 128].float().cpu(), ref) and float(y[:, 8 * 6 * 128:].abs().max()) == 0.0


def test_discovery_fitness_with_resnet_assessor_vs_reference(engine, golden_resnet):
    """Whole fitness (G -> per-image rescale -> ResNet-50 -> softmax pick -> clamp/affine) vs the reference's own
    `Discovery.fitness` outputs: |dfit| <= 2e-2 (probabilities through 56 bf16 layers)."""
    g = golden_resnet
    gen = port.Generator(100, 1, 16).eval()
    gen.load_state_dict({k[len("fit/gen/"):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("fit/gen/")})
    m = _model(g, "r50_c8")
    pos = torch.from_numpy(g["fit/pos"]).float().cuda()
    P = pos.shape[0]
    gnet = engine.GeneratorNet(gen.state_dict(), P)
    anet = engine.AssessorNet(m.state_dict(), P)
    for mode in ("optimize_out_training", "optimize_in_training"):
        sw = engine.SwarmEngine(P, 100, generator=gnet, fitness=engine.ResnetFitness(anet, 2, mode), seed=1)
        sw.init_particles(pos=pos, vel=torch.zeros_like(pos))
        sw.step()
        torch.cuda.synchronize()
        assert np.abs(sw.fit.cpu().numpy() - g[f"fit/{mode}"]).max() <= 2e-2


def test_dropin_discovery_with_mirror_resnet(engine, golden_resnet):
    """The reference's call sequence (src/training/pso_discovery.py:177-219): get_cnn('ResNet50', ...), load_state_dict,
    Discovery(...).fitness(dim_space, pos) per particle -> the reference's own fitness values, and a whole Swarm."""
    import tempfile
    from src.pso import util_cnn, util_discovery, util_pso
    from src.utils import util_dcgan
    g = golden_resnet
    classes = [0, 2, 3, 4, 6, 7, 8, 9]
    ref_cnn = _model(g, "r50_c8")
    cnn = util_cnn.get_cnn("ResNet50", 1, iid_classes=classes, n_class=8).cuda()
    assert list(cnn.state_dict().keys()) == list(ref_cnn.state_dict().keys())
    cnn.load_state_dict(ref_cnn.state_dict())
    gen = util_dcgan.Generator(100, 1, 16).cuda()
    gen.load_state_dict({k[len("fit/gen/"):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("fit/gen/")})
    x = torch.from_numpy(g["x"]).cuda()
    assert np.abs(cnn(x).cpu().numpy() - g["r50_c8/logits"]).max() <= 0.03 * np.abs(g["r50_c8/logits"]).max()
    for mode in ("optimize_out_training", "optimize_in_training"):
        d = util_discovery.Discovery(iid_class=3, model_gan=gen, model_cnn=cnn, device="cuda", control_pso_fitness=mode,
                                     iid_classes=classes)
        fit = np.concatenate([d.fitness(100, g["fit/pos"][i])[0] for i in range(g["fit/pos"].shape[0])])
        assert fit.dtype == np.float32 and np.abs(fit - g[f"fit/{mode}"]).max() <= 2e-2
    with tempfile.TemporaryDirectory() as td:
        sw = util_pso.Swarm(td, d, num_particles=64, n_iterations=4, dim_space=100, device="cuda", tolerance=1e-5,
                            w_inertia=0.73, w_cogn=1.496, w_soci=1.496)
        history, _, _, it = sw.optimize(schedule_inertia=False, early_stopping=False)
    assert it == 5 and 0.1 <= float(np.asarray(sw.g_best_val[-1]).reshape(-1)[0]) <= 1.1
    assert all(b <= a for a, b in zip(sw.g_best_val, sw.g_best_val[1:]))


def test_full_size_assessor_size_independent_properties(engine, golden_resnet):
    """ResNet-50 assessor at the bench size (65 536 images): an image's logits do not depend on its batch (whole batch ==
    two halves == reversed order, bit for bit); the golden images planted in the batch give the golden logits; 24 random
    rows agree with the fp32 oracle."""
    P = 65536
    m = _model(golden_resnet, "r50_c8")
    gen = torch.Generator(device="cuda").manual_seed(4)
    base = torch.rand(P, 784, generator=gen, device="cuda")
    # structured images: low-frequency fields, so that logits differ between rows
    coarse = torch.rand(P, 1, 4, 4, generator=gen, device="cuda")
    x = (0.3 * base + torch.nn.functional.interpolate(coarse, size=28, mode="bilinear").reshape(P, 784)).contiguous()
    gx = torch.from_numpy(golden_resnet["x"]).reshape(-1, 784).cuda()
    x[1000:1000 + gx.shape[0]] = gx
    x[P - gx.shape[0]:] = gx
    whole = engine.AssessorNet(m.state_dict(), P).forward(x).clone()
    assert bool(torch.isfinite(whole).all())
    half = engine.AssessorNet(m.state_dict(), P // 2)
    a = half.forward(x[:P // 2].contiguous()).clone()
    b = half.forward(x[P // 2:].contiguous()).clone()
    assert torch.equal(torch.cat([a, b]), whole)
    del half
    rev = engine.AssessorNet(m.state_dict(), P).forward(x.flip(0).contiguous())
    assert torch.equal(rev.flip(0), whole)
    ref = golden_resnet["r50_c8/logits"]
    for s in (1000, P - gx.shape[0]):
        assert np.abs(whole[s:s + gx.shape[0]].cpu().numpy() - ref).max() <= 0.03 * np.abs(ref).max()
    idx = torch.randperm(P, generator=torch.Generator().manual_seed(2))[:24]
    with torch.no_grad():
        oracle = m(x[idx.cuda()].cpu().view(-1, 1, 28, 28)).numpy()
    assert np.abs(whole[idx.cuda()].cpu().numpy() - oracle).max() <= 0.03 * np.abs(oracle).max()
