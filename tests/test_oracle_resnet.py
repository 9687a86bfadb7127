"""CPU: the oracle's restatement of the reference assessor (`oracle.port.Resnet`, src/pso/util_cnn.py:81-190)
against outputs of the reference's own class (tests/golden/ref_resnet.npz, oracle/gen_golden_resnet.py)."""
import numpy as np
import pytest
import torch

from oracle import port

CASES = ["r50_c8", "r50_c2", "r1111_c8", "r2121_c3"]


def build(golden_resnet, name):
    *layers, n_class, seed = (int(v) for v in golden_resnet[f"{name}/meta"])
    m = port.Resnet(layers, 1, list(range(n_class)) if n_class > 2 else [0, 1], n_class).eval()
    port.seeded_fill(m, seed)
    return m


@pytest.mark.parametrize("name", CASES)
def test_port_resnet_matches_reference_outputs(golden_resnet, name):
    m = build(golden_resnet, name)
    sd = m.state_dict()
    keys = sorted(k for k in sd if not k.endswith("num_batches_tracked"))
    # same parameters as the reference model had (state_dict keys and shapes agree, or the sums would not)
    assert np.array_equal(np.array([float(sd[k].double().sum()) for k in keys]), golden_resnet[f"{name}/sum"])
    assert np.array_equal(np.array([float(sd[k].double().abs().sum()) for k in keys]), golden_resnet[f"{name}/abssum"])
    x = torch.from_numpy(golden_resnet["x"])
    with torch.no_grad():
        np.testing.assert_allclose(m.forward_avgpool(x).numpy(), golden_resnet[f"{name}/features"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(m(x).numpy(), golden_resnet[f"{name}/logits"], rtol=1e-5, atol=1e-5)


def test_port_fitness_with_resnet_assessor_matches_reference_discovery(golden_resnet):
    g = golden_resnet
    gen = port.Generator(100, 1, 16).eval()
    gen.load_state_dict({k[len("fit/gen/"):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("fit/gen/")})
    m = build(g, "r50_c8")
    for mode in ("optimize_out_training", "optimize_in_training"):
        fit = np.concatenate([port.fitness_assessor(gen, m, g["fit/pos"][i], 100, mode=mode, class_index=2)[0]
                              for i in range(g["fit/pos"].shape[0])])
        np.testing.assert_allclose(fit, g[f"fit/{mode}"], rtol=1e-5, atol=1e-6)


def test_port_bn_free_vq_decoder_matches_reference(golden_resnet):
    """`VectorQuantizedVAE_MNIST.decoder` (src/inverter/utils_vq_vae/util_model.py:237-244) restated, and its lowering to
    engine tables (numpy emulation, bf16 operands)."""
    from psob import nets
    g = golden_resnet
    m = port.VQDecoderMNIST(1, 100, 16).eval()
    sd = {k[len("vqmnist/"):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("vqmnist/decoder.")}
    m.load_state_dict(sd)
    with torch.no_grad():
        out = m(torch.from_numpy(g["vqmnist/z"])).numpy()
    np.testing.assert_allclose(out, g["vqmnist/out"], rtol=1e-5, atol=1e-6)
    progs = nets.generator_programs({k: v.numpy() for k, v in sd.items()}, prefix="decoder")
    z = np.zeros((6, progs[0].in_pitch), np.float32)
    from psob import lowering as L
    z[:, :100] = L.to_bf16_f32(g["vqmnist/z"].reshape(6, 100))
    _, img = nets.emulate_chain(progs, z)
    assert np.abs(img[:, :784] - g["vqmnist/out"].reshape(6, 784)).max() <= 2e-2
