"""TEST INFRASTRUCTURE -- generate `tests/golden/ref_vq.npz` from the LIVE reference VQ-VAE classes.

    python oracle/gen_golden_vq.py          (build container only: needs /root/reference)

Unmodified `VectorQuantizedVAE` (src/inverter/utils_vq_vae/util_model.py:179-222: the ResBlockBatchNorm decoder on the 7x7
latent grid, :194-203, block :151-164), `VQEmbedding` + `pso_weights` (:49-54,125-131: PSO positions -> nn.Embedding.weight)
and `vq` (util_function.py:4-23) on seeded inputs:

  vqgrid/*      decoder state_dict (BN statistics randomised, eval mode), codebook set through pso_weights from a float32
                DataFrame like training/vq_vae.py:35-57 builds it, decoder(z) for z [6, 100, 7, 7], decode(latents) for index
                maps [4, 7, 7]
  vqsnap/*      the VQ-snapped variant of BASELINE config 3: `VectorQuantizedVAE_GAN` (features 16) with a pso_weights codebook;
                positions x [12, 100] -> vq -> codebook rows -> decoder
"""
import contextlib
import io
import os
import sys

import numpy as np
import pandas as pd
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import port, ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "ref_vq.npz")


def put_sd(g, prefix, sd, keep):
    for k, v in sd.items():
        if k.startswith(keep) and not k.endswith("num_batches_tracked"):
            g[f"{prefix}/{k}"] = v.detach().numpy().copy()


def main():
    ref = ref_loader.load()
    V, F = ref.vq_model, ref.vq_function
    rng = np.random.RandomState(2024)
    g = {}
    # ---- VectorQuantizedVAE: ResBlockBN decoder on the 7x7 grid, codebook from PSO positions -----------------
    data = pd.DataFrame((0.5 * rng.randn(256, 100)).astype("float32"))
    torch.manual_seed(51)
    with contextlib.redirect_stdout(io.StringIO()):
        m = V.VectorQuantizedVAE(1, 100, 256, data).eval()
    assert m.codebook.embedding.weight.dtype == torch.float32
    assert np.array_equal(m.codebook.embedding.weight.detach().numpy(), data.values)          # pso_weights (:49-54)
    port.randomize_bn_stats(m, torch.Generator().manual_seed(52))
    z = torch.randn(6, 100, 7, 7, generator=torch.Generator().manual_seed(53))
    latents = torch.randint(0, 256, (4, 7, 7), generator=torch.Generator().manual_seed(54))
    with torch.no_grad():
        g["vqgrid/z"] = z.numpy().copy()
        g["vqgrid/out"] = m.decoder(z.clone()).numpy()            # the block's first ReLU is in place: hand it a copy
        g["vqgrid/latents"] = latents.numpy()
        g["vqgrid/decode_out"] = m.decode(latents).numpy()
    g["vqgrid/codebook"] = data.values.copy()
    put_sd(g, "vqgrid", m.state_dict(), "decoder.")
    # ---- VQ-snapped config 3: positions -> nearest code -> VectorQuantizedVAE_GAN.decoder --------------------
    data2 = pd.DataFrame(rng.randn(64, 100).astype("float32"))
    torch.manual_seed(61)
    with contextlib.redirect_stdout(io.StringIO()):
        m2 = V.VectorQuantizedVAE_GAN(1, 100, 64, data2, features_g=16, features_d=16).eval()
    port.randomize_bn_stats(m2, torch.Generator().manual_seed(62))
    x = torch.from_numpy((rng.randn(12, 100) * 1.2).astype("float32"))
    with torch.no_grad():
        idx = F.vq(x, m2.codebook.embedding.weight)
        snapped = m2.codebook.embedding(idx)
        g["vqsnap/x"] = x.numpy()
        g["vqsnap/idx"] = idx.numpy()
        g["vqsnap/out"] = m2.decoder(snapped.view(12, 100, 1, 1)).numpy()
    g["vqsnap/codebook"] = data2.values.copy()
    put_sd(g, "vqsnap", m2.state_dict(), "decoder.")
    np.savez_compressed(OUT, **g)
    print(f"wrote {OUT}: {len(g)} arrays, {os.path.getsize(OUT) / 1e6:.2f} MB")
    print("vqgrid/out range", g["vqgrid/out"].min(), g["vqgrid/out"].max(), "std", g["vqgrid/out"].std())


if __name__ == "__main__":
    main()
