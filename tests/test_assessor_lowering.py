"""CPU: lowering of the reference's ResNet assessor (src/pso/util_cnn.py:81-190) to pixel-GEMM tables, checked through
the numpy emulation of the engine against the fp32 torch restatement and the reference's golden logits."""
import numpy as np
import pytest
import torch

from oracle import port
from psob import lowering as L, nets


def _model(golden_resnet, name):
    *layers, n_class, seed = (int(v) for v in golden_resnet[f"{name}/meta"])
    m = port.Resnet(layers, 1, list(range(n_class)) if n_class > 2 else [0, 1], n_class).eval()
    port.seeded_fill(m, seed)
    return m


@pytest.mark.parametrize("mode", ["mma", "epilogue"])
@pytest.mark.parametrize("name", ["r1111_c8", "r2121_c3"])
def test_assessor_tables_are_the_reference_arithmetic_in_fp32(golden_resnet, name, mode, monkeypatch):
    """Without bf16 rounding the tables reproduce the fp32 network to summation-order noise: geometry, BN folding,
    residual wiring, max pool and border handling are exact."""
    monkeypatch.setattr(L, "ROUND_WEIGHTS", False)
    m = _model(golden_resnet, name)
    ops, bufs = nets.assessor_programs(m.state_dict(), shortcut=mode)
    x = golden_resnet["x"]
    got = nets.emulate_assessor(ops, bufs, x.reshape(x.shape[0], -1), round_out=False)
    ref = golden_resnet[f"{name}/logits"]
    assert np.abs(got - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())


def test_resnet50_tables_bf16_vs_reference_logits(golden_resnet):
    """bf16 weights and activations, fp32 accumulation (what the engine computes): logits within 3 % of the logit
    range of the fp32 reference over 53 layers; exact MAC count of the border-aware network."""
    m = _model(golden_resnet, "r50_c8")
    ops, bufs = nets.assessor_programs(m.state_dict())
    x = golden_resnet["x"]
    got = nets.emulate_assessor(ops, bufs, x.reshape(x.shape[0], -1))
    ref = golden_resnet["r50_c8/logits"]
    assert np.abs(got - ref).max() <= 0.03 * np.abs(ref).max()
    sm = lambda v: torch.softmax(torch.from_numpy(v), 1).numpy()
    assert np.abs(sm(got) - sm(ref)).max() <= 3e-2
    progs = [o.prog for o in ops if o.prog is not None]
    assert len(progs) == 1 + 16 * 3 + 1           # projection shortcuts are folded into the third conv
    # border-aware MACs counted independently with torch: conv(ones, ones) counts the taps inside the map
    macs = 0
    h = 28
    def conv_macs(cin, cout, k, s, p, h):
        taps = torch.nn.functional.conv2d(torch.ones(1, 1, h, h), torch.ones(1, 1, k, k), stride=s, padding=p)
        return int(taps.sum().item()) * cin * cout, taps.shape[-1]
    n, h = conv_macs(1, 64, 7, 2, 3, 28); macs += n
    h = (h + 2 - 3) // 2 + 1
    cin = 64
    for nb, mid, stride in zip((3, 4, 6, 3), (64, 128, 256, 512), (1, 2, 2, 2)):
        for b in range(nb):
            s = stride if b == 0 else 1
            macs += conv_macs(cin, mid, 1, 1, 0, h)[0]
            n, ho = conv_macs(mid, mid, 3, s, 1, h); macs += n
            macs += conv_macs(mid, 4 * mid, 1, 1, 0, ho)[0]
            if b == 0:
                macs += conv_macs(cin, 4 * mid, 1, s, 0, h)[0]
            cin, h = 4 * mid, ho
    macs += 2048 * 8
    assert sum(p.macs_exact for p in progs) == macs
    assert all((p.meta.get("shortcut") is not None) == p.name.endswith(".c3") for p in progs)
    ops_e, _ = nets.assessor_programs(m.state_dict(), shortcut="epilogue")
    progs_e = [o.prog for o in ops_e if o.prog is not None]
    assert len(progs_e) == 1 + 16 * 3 + 4 + 1 and sum(p.macs_exact for p in progs_e) == macs
    assert all(p.meta.get("residual", False) == p.name.endswith(".c3") for p in progs_e)


def test_conv2d_lowering_matches_torch_on_odd_shapes(monkeypatch):
    monkeypatch.setattr(L, "ROUND_WEIGHTS", False)
    g = torch.Generator().manual_seed(5)
    for (cin, cout, k, s, p, h) in [(64, 64, 3, 1, 1, 7), (128, 128, 3, 2, 1, 7), (256, 512, 1, 2, 0, 7), (64, 256, 1, 1, 0, 5),
                                    (512, 512, 3, 1, 1, 1), (128, 192, 3, 2, 1, 4)]:
        w = torch.randn(cout, cin, k, k, generator=g) * 0.05
        b = torch.randn(cout, generator=g)
        x = torch.randn(3, cin, h, h, generator=g)
        prog = L.lower_conv2d("t", w.numpy(), b.numpy(), L.ACT_RELU, h, h, s, p)
        ref = torch.relu(torch.nn.functional.conv2d(x, w, b, stride=s, padding=p))
        y, _ = L.emulate(prog, x.permute(0, 2, 3, 1).reshape(3, -1).numpy(), round_out=False)
        ho = ref.shape[-1]
        got = y.reshape(3, ho, ho, L.cpitch(cout))[..., :cout].transpose(0, 3, 1, 2)
        assert np.abs(got - ref.numpy()).max() <= 1e-4
        taps = torch.nn.functional.conv2d(torch.ones(1, 1, h, h), torch.ones(1, 1, k, k), stride=s, padding=p).sum().item()
        assert prog.macs_exact == int(taps) * cin * cout
