"""Generator / Discriminator with the reference's constructor signatures and state_dict keys
(src/utils/util_dcgan.py:103-149), whose forward runs on the B200 pixel-GEMM engine.

The modules own ordinary fp32 parameters so `load_state_dict(best_g.tar['model_state_dict'])`
works unchanged (src/training/pso_discovery.py:195-198); `forward` folds BatchNorm (eval
semantics — the only mode the PSO path uses, src/pso/util_discovery.py:45), lowers the stack to
bf16 tcgen05 programs once per (batch size, parameter version) and executes them.  There is no
CPU path: tensors must live on a CUDA device."""
from __future__ import annotations

import torch
import torch.nn as nn

from psob import engine as _engine
from psob import _ffi


def initialize_weights(model):
    """N(0, 0.02) on conv / conv-transpose / BN weights (src/utils/util_dcgan.py:45-48)."""
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d, nn.BatchNorm2d)):
            nn.init.normal_(m.weight.data, 0.0, 0.02)


class _EngineModule(nn.Module):
    _net_cls = None
    _prefix = None

    def _version(self):
        return tuple(int(t._version) for t in self.state_dict().values())

    def engine_net(self, batch: int):
        key = (batch, self._version(), str(next(self.parameters()).device))
        cache = self.__dict__.setdefault("_nets", {})
        if key not in cache:
            cache.clear()
            cache[key] = self._net_cls(self.state_dict(), batch, prefix=self._prefix, device=next(self.parameters()).device)
        return cache[key]

    def _check(self, x):
        if not x.is_cuda:
            raise _ffi.PsobError(f"{type(self).__name__}.forward needs CUDA tensors (B200); there is no CPU fallback")


def _up(cin, cout, k, s, p):
    return nn.Sequential(nn.ConvTranspose2d(cin, cout, k, s, p), nn.BatchNorm2d(cout), nn.ReLU())


class Generator(_EngineModule):
    _net_cls = _engine.GeneratorNet
    _prefix = "gen"

    def __init__(self, z_dim, channels_img, features_g):
        super().__init__()
        if channels_img != 1:
            raise _ffi.PsobError("the B200 generator path implements single-channel images (the reference's PSO configs)")
        self.z_dim = z_dim
        self.gen = nn.Sequential(
            self._block(z_dim, features_g * 2, 7, 1, 0),
            self._block(features_g * 2, features_g, 4, 2, 1),
            nn.ConvTranspose2d(features_g, channels_img, 4, 2, 1),
            nn.Tanh(),
        )

    def _block(self, in_channels, out_channels, kernel_size, stride, padding):
        return _up(in_channels, out_channels, kernel_size, stride, padding)

    @torch.no_grad()
    def forward(self, x):
        self._check(x)
        n = x.shape[0]
        net = self.engine_net(n)
        z = torch.zeros(n, net.in_pitch, dtype=torch.bfloat16, device=x.device)
        _engine.pack_latent(x.reshape(n, -1).float().contiguous(), z)
        _, img = net.forward(z)
        return img.view(n, 1, 28, 28).clone()


class Discriminator(_EngineModule):
    _net_cls = _engine.DiscriminatorNet
    _prefix = "disc"

    def __init__(self, channels_img, features_d):
        super().__init__()
        if channels_img != 1:
            raise _ffi.PsobError("the B200 discriminator path implements single-channel images")
        self.disc = nn.Sequential(
            nn.Conv2d(channels_img, features_d, 4, 2, 1),
            nn.LeakyReLU(0.2),
            self._block(features_d, features_d * 2, 4, 2, 1),
            nn.Conv2d(features_d * 2, 1, 7, 2, 0),
            nn.Sigmoid(),
        )

    def _block(self, in_channels, out_channels, kernel_size, stride, padding):
        return nn.Sequential(nn.Conv2d(in_channels, out_channels, kernel_size, stride, padding), nn.LeakyReLU(0.2))

    @torch.no_grad()
    def forward(self, x):
        self._check(x)
        n = x.shape[0]
        net = self.engine_net(n)
        xp = torch.empty(n, 31 * 32, dtype=torch.bfloat16, device=x.device)
        _engine.pack_image(x.reshape(n, -1).float().contiguous(), xp)
        _, score = net.forward(xp)
        return score.view(n, 1, 1, 1).clone()


class Encoder(_EngineModule):
    """Inverter encoder (src/inverter/utils_ae/util_inverter.py:164-184)."""
    _net_cls = _engine.EncoderNet
    _prefix = "enc"

    def __init__(self, enc_dim, channels_img, features_e=64):
        super().__init__()
        self.enc_dim = enc_dim
        self.enc = nn.Sequential(
            nn.Conv2d(channels_img, features_e, 4, 2, 1),
            nn.LeakyReLU(0.2),
            nn.Sequential(nn.Conv2d(features_e, features_e * 2, 4, 2, 1), nn.LeakyReLU(0.2)),
            nn.Conv2d(features_e * 2, enc_dim, 7, 2, 0),
        )

    @torch.no_grad()
    def forward(self, x):
        self._check(x)
        n = x.shape[0]
        net = self.engine_net(n)
        xp = torch.empty(n, 31 * 32, dtype=torch.bfloat16, device=x.device)
        _engine.pack_image(x.reshape(n, -1).float().contiguous(), xp)
        _, lat = net.forward(xp)
        return lat.view(n, self.enc_dim, 1, 1).clone()
