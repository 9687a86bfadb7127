"""TEST INFRASTRUCTURE -- generate `tests/golden/ref_resnet.npz` from the LIVE reference assessor.

    python oracle/gen_golden_resnet.py          (build container only: needs /root/reference)

The reference's `Resnet` / `resnet_block` (src/pso/util_cnn.py:81-190) are imported unmodified, their
parameters filled by `oracle.port.seeded_fill` (seeded, order-independent -- a ResNet-50 state_dict is
94 MB, so the fixture stores the seed, the inputs and the reference's outputs, plus per-tensor checksums
that prove the port and the tests rebuilt the same parameters).  Also the real `Discovery.fitness`
(src/pso/util_discovery.py:52-82) with this assessor behind a seeded generator.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import port, ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "ref_resnet.npz")
CASES = {"r50_c8": ([3, 4, 6, 3], 8, 11), "r50_c2": ([3, 4, 6, 3], 2, 12), "r1111_c8": ([1, 1, 1, 1], 8, 13), "r2121_c3": ([2, 1, 2, 1], 3, 14)}


def checksums(model):
    sd = model.state_dict()
    keys = sorted(k for k in sd if not k.endswith("num_batches_tracked"))
    return np.array([float(sd[k].double().sum()) for k in keys]), np.array([float(sd[k].double().abs().sum()) for k in keys])


def main():
    ref = ref_loader.load()
    C = ref.util_cnn
    g = {}
    tg = torch.Generator().manual_seed(77)
    # inputs: min-max rescaled images in [0, 1] like Discovery feeds the assessor (util_discovery.py:61-66)
    yy, xx = torch.meshgrid(torch.arange(28.0), torch.arange(28.0), indexing="ij")
    imgs = [
        torch.nn.functional.interpolate(torch.rand(1, 1, 4, 4, generator=tg), size=28, mode="bilinear")[0, 0],   # smooth field
        (xx + 2 * yy) / 81.0,                                                                                      # ramp
        ((xx // 4 + yy // 4) % 2),                                                                                 # checkerboard
        (torch.rand(28, 28, generator=tg) > 0.97).float(),                                                         # sparse dots
        (((xx - 13.5) ** 2 + (yy - 11.0) ** 2) < 60).float() + 0.2 * torch.rand(28, 28, generator=tg),            # disc + noise
        torch.rand(28, 28, generator=tg),                                                                          # white noise
        torch.exp(-((xx - 8) ** 2 + (yy - 19) ** 2) / 18.0),                                                       # blob
        (torch.abs(xx - yy) < 2).float() + (torch.abs(xx + yy - 27) < 2).float(),                                  # cross
    ]
    x = port.rescale_minmax_per_image(torch.stack(imgs).unsqueeze(1).float())
    g["x"] = x.numpy()
    for name, (layers, n_class, seed) in CASES.items():
        classes = list(range(n_class)) if n_class > 2 else [0, 1]
        torch.manual_seed(seed)
        m = C.Resnet(C.resnet_block, layers, 1, iid_classes=classes, n_class=n_class).eval()
        port.seeded_fill(m, seed)
        with torch.no_grad():
            g[f"{name}/logits"] = m(x).numpy()
            g[f"{name}/features"] = m.forward_avgpool(x).numpy()
        g[f"{name}/sum"], g[f"{name}/abssum"] = checksums(m)
        g[f"{name}/meta"] = np.array(layers + [n_class, seed])
        print(name, "logits", np.round(g[f"{name}/logits"][:2], 3))
    # the real Discovery.fitness with the ResNet-50 assessor and a seeded reference Generator (batch 1, per particle)
    torch.manual_seed(5)
    gen = ref.util_dcgan.Generator(100, 1, 16).eval()
    ref.util_dcgan.initialize_weights(gen)
    port.randomize_bn_stats(gen, torch.Generator().manual_seed(6))
    m = C.Resnet(C.resnet_block, [3, 4, 6, 3], 1, iid_classes=list(range(8)), n_class=8).eval()
    port.seeded_fill(m, 11)
    pos = torch.randn(5, 100, generator=tg).double().numpy()
    classes = [0, 2, 3, 4, 6, 7, 8, 9]
    for mode in ("optimize_out_training", "optimize_in_training"):
        d = ref.util_discovery.Discovery(iid_class=3, model_gan=gen, model_cnn=m, device="cpu", control_pso_fitness=mode,
                                         iid_classes=classes, obj_fun_threshold=0.0)
        m.class_to_idx = {c: i for i, c in enumerate(classes)}
        out = [d.fitness(100, pos[i]) for i in range(pos.shape[0])]
        g[f"fit/{mode}"] = np.concatenate([o[0] for o in out])
        g["fit/img_rescaled"] = np.concatenate([o[1].numpy() for o in out])
    g["fit/pos"] = pos
    for k, v in gen.state_dict().items():
        g[f"fit/gen/{k}"] = v.numpy()
    print("fitness", g["fit/optimize_out_training"], g["fit/optimize_in_training"])
    # the BatchNorm-free VQ-VAE decoder (VectorQuantizedVAE_MNIST.decoder, util_model.py:237-244), small widths, with
    # non-zero biases so that the bias path is exercised
    torch.manual_seed(21)
    vq = ref.vq_model.VectorQuantizedVAE_MNIST(1, embedded_dim=100, num_embedding=32, data_pso=None, num_hiddens=16).eval()
    with torch.no_grad():
        for m_ in vq.decoder:
            if hasattr(m_, "bias") and m_.bias is not None:
                m_.bias.copy_(torch.randn(m_.bias.shape, generator=tg) * 0.1)
        zq = torch.randn(6, 100, 1, 1, generator=tg)
        g["vqmnist/z"] = zq.numpy()
        g["vqmnist/out"] = vq.decoder(zq).numpy()
    for k, v in vq.state_dict().items():
        if k.startswith("decoder."):
            g[f"vqmnist/{k}"] = v.numpy()
    np.savez_compressed(OUT, **g)
    print("wrote", OUT, os.path.getsize(OUT) // 1024, "KiB,", len(g), "arrays")


if __name__ == "__main__":
    main()
