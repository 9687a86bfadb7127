"""VQ-VAE pieces on the swarm-evaluation path, with the reference's names and state_dict keys
(src/inverter/utils_vq_vae/util_model.py): `pso_weights` (:49-54), `VQEmbedding` (:125-148, forward only),
`ResBlockBatchNorm` (:151-164), `VectorQuantizedVAE` (:179-222: `decode` / `decoder` on the B200 engine) and
`VectorQuantizedVAE_GAN` (:268-300: `decoder` == the DCGAN generator stack).

In scope: the decoders, the codebook lookup (`vq`), `pso_weights`.  Out of scope, and raising
if called: encoders' training-time use, the straight-through estimator and every loss (training code).
The modules own ordinary fp32 parameters, so reference checkpoints load with `load_state_dict` unchanged."""
from __future__ import annotations

import torch
import torch.nn as nn

from psob import engine as _engine
from psob import _ffi
from src.inverter.utils_vq_vae import util_function
from src.utils import util_dcgan


def weights_init(m):
    """xavier-uniform conv weights, zero biases (src/inverter/utils_vq_vae/util_model.py:39-46)."""
    classname = m.__class__.__name__
    if classname.find("Conv") != -1:
        try:
            nn.init.xavier_uniform_(m.weight.data)
            m.bias.data.fill_(0)
        except AttributeError:
            print("Skipping initialization of ", classname)


@torch.no_grad()
def pso_weights(model, data):
    """PSO positions -> codebook: every nn.Embedding of `model` gets `data.values` as its weight (:49-54).  `data` is the float32
    DataFrame of the particles' final positions (one row per particle, training/vq_vae.py:35-57)."""
    data_tensor = torch.tensor(data.values)
    for m in model.modules():
        if isinstance(m, nn.Embedding):
            m.weight.data = data_tensor.to(m.weight.device)


class VQEmbedding(nn.Module):
    def __init__(self, K, D, data_pso):
        super().__init__()
        self.embedding = nn.Embedding(K, D)
        if data_pso is not None:
            pso_weights(self.embedding, data_pso)
        else:
            self.embedding.weight.data.uniform_(-1. / K, 1. / K)

    @torch.no_grad()
    def forward(self, z_e_x):
        """[B, D, H, W] -> nearest-code indices [B, H, W] (psob_vq_nearest)."""
        z_e_x_ = z_e_x.permute(0, 2, 3, 1).contiguous()
        return util_function.vq(z_e_x_, self.embedding.weight)

    def straight_through(self, z_e_x):
        raise _ffi.PsobError("VQEmbedding.straight_through is training code (vq_st backward): out of the swarm-evaluation scope")


class ResBlockBatchNorm(nn.Module):
    """Parameter holder with the reference's keys (`block.{1,2,4,5}.*`); evaluated by VQGridDecoderNet."""

    def __init__(self, dim):
        super().__init__()
        self.block = nn.Sequential(nn.ReLU(True), nn.Conv2d(dim, dim, 3, 1, 1), nn.BatchNorm2d(dim), nn.ReLU(True),
                                   nn.Conv2d(dim, dim, 1), nn.BatchNorm2d(dim))


class _GridDecoder(nn.Sequential):
    """`VectorQuantizedVAE.decoder`: an nn.Sequential (same keys as the reference's) whose forward runs on the engine."""

    def _version(self):
        return tuple(int(t._version) for t in self.state_dict().values())

    def engine_net(self, batch: int):
        dev = next(self.parameters()).device
        key = (batch, self._version(), str(dev))
        cache = self.__dict__.setdefault("_nets", {})
        if key not in cache:
            cache.clear()
            sd = {"decoder." + k: v for k, v in self.state_dict().items()}
            cache[key] = _engine.VQGridDecoderNet(sd, batch, prefix="decoder", device=dev)
        return cache[key]

    @torch.no_grad()
    def forward(self, x):
        if not x.is_cuda:
            raise _ffi.PsobError("VectorQuantizedVAE.decoder needs CUDA tensors (B200); there is no CPU fallback")
        n = x.shape[0]
        img = self.engine_net(n).forward(x.float().contiguous())
        return img.view(n, 1, 28, 28).clone()


class VectorQuantizedVAE(nn.Module):
    def __init__(self, channels_img, embedded_dim=64, num_embedding=512, data_pso=None):
        super().__init__()
        if channels_img != 1:
            raise _ffi.PsobError("the B200 decoder path implements single-channel images")
        dim = embedded_dim
        self.encoder = nn.Sequential(nn.Conv2d(channels_img, dim, 4, 2, 1), nn.BatchNorm2d(dim), nn.ReLU(True), nn.Conv2d(dim, dim, 4, 2, 1),
                                     ResBlockBatchNorm(dim), ResBlockBatchNorm(dim))
        self.codebook = VQEmbedding(num_embedding, embedded_dim, data_pso=data_pso)
        self.decoder = _GridDecoder(ResBlockBatchNorm(dim), ResBlockBatchNorm(dim), nn.ReLU(True), nn.ConvTranspose2d(dim, dim, 4, 2, 1),
                                    nn.BatchNorm2d(dim), nn.ReLU(True), nn.ConvTranspose2d(dim, channels_img, 4, 2, 1), nn.Tanh())
        self.apply(weights_init)

    @torch.no_grad()
    def decode(self, latents):
        """index map [B, 7, 7] -> images [B, 1, 28, 28] (:213-216)."""
        z_q_x = self.codebook.embedding(latents).permute(0, 3, 1, 2)
        return self.decoder(z_q_x)

    def encode(self, x):
        raise _ffi.PsobError("VectorQuantizedVAE.encode: the encoder is training-side code, out of the swarm-evaluation scope")

    def forward(self, x):
        raise _ffi.PsobError("VectorQuantizedVAE.forward is the training pass (straight-through estimator): out of scope")


class VectorQuantizedVAE_GAN(nn.Module):
    """`decoder` is the DCGAN generator stack (:283-289) -- `src.utils.util_dcgan.Generator` under the `decoder.*` keys --
    and `decode_snapped` evaluates positions at their nearest code (the VQ-snapped variant of BASELINE config 3)."""

    def __init__(self, channels_img, embedded_dim=64, num_embedding=512, data_pso=None, features_g=64, features_d=64):
        super().__init__()
        # parameter holder with the reference's keys (checkpoints load unchanged); evaluating it is training-side code
        self.encoder = nn.Sequential(
            nn.Conv2d(channels_img, features_d, 4, 2, 1), nn.LeakyReLU(0.2),
            nn.Sequential(nn.Conv2d(features_d, features_d * 2, 4, 2, 1), nn.BatchNorm2d(features_d * 2), nn.LeakyReLU(0.2)),
            nn.Conv2d(features_d * 2, embedded_dim, 7, 2, 0))
        self.codebook = VQEmbedding(K=num_embedding, D=embedded_dim, data_pso=data_pso)
        g = util_dcgan.Generator(embedded_dim, channels_img, features_g)
        self._gen = [g]                          # not a registered sub-module: `decoder` owns the parameters
        self.decoder = g.gen
        self.apply(weights_init)

    def encode(self, x):
        raise _ffi.PsobError("VectorQuantizedVAE_GAN.encode: the encoder is training-side code, out of the swarm-evaluation scope")

    @torch.no_grad()
    def decode(self, latents):
        """index map [B, 1, 1] -> images (:309-312)."""
        z_q_x = self.codebook.embedding(latents).permute(0, 3, 1, 2)
        return self._generator()(z_q_x.contiguous())

    def _generator(self):
        return self._gen[0]

    @torch.no_grad()
    def decode_snapped(self, x):
        """x [N, D] -> (indices [N], images [N, 1, 28, 28]): `vq` then the decoder on the selected codebook rows."""
        idx = util_function.vq(x.float(), self.codebook.embedding.weight)
        z = self.codebook.embedding(idx)
        return idx, self._generator()(z.view(x.shape[0], -1, 1, 1))

    def forward(self, x):
        raise _ffi.PsobError("VectorQuantizedVAE_GAN.forward is the training pass: out of scope")
