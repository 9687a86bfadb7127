"""TEST INFRASTRUCTURE — CPU restatement ("port") of the reference's swarm-evaluation path.

This file is the *oracle*: a plain numpy / torch-fp32-on-CPU restatement of what
ltronchin/Gan-discovery-PSO computes on the hot path.  It is pinned against the
live reference classes by `oracle/gen_golden.py` (run in the build container,
where `/root/reference` exists) and the committed fixtures in `tests/golden/`.
It is imported only by `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py`.  The product
(`gan-discovery-pso_b200/`) never imports it and has no CPU fallback.

Citations are `file:line` relative to the reference root.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# Models (fp32 torch, CPU).  Same constructor signatures and state_dict keys as the
# reference so checkpoints / golden weights load into either.
# --------------------------------------------------------------------------------------


def _up(cin: int, cout: int, k: int, s: int, p: int) -> nn.Sequential:
    # ConvTranspose2d -> BatchNorm2d -> ReLU   (src/utils/util_dcgan.py:141-146)
    return nn.Sequential(nn.ConvTranspose2d(cin, cout, k, s, p), nn.BatchNorm2d(cout), nn.ReLU())


def _down(cin: int, cout: int, k: int, s: int, p: int) -> nn.Sequential:
    # Conv2d -> LeakyReLU(0.2), no BN          (src/utils/util_dcgan.py:116-121)
    return nn.Sequential(nn.Conv2d(cin, cout, k, s, p), nn.LeakyReLU(0.2))


class Generator(nn.Module):
    """DCGAN generator z[N,z,1,1] -> [N,C,28,28]  (src/utils/util_dcgan.py:128-149;
    verbatim twins at src/inverter/utils_ae/util_inverter.py:119-140 and
    src/inverter/utils_vq_vae/util_model.py:57-78)."""

    def __init__(self, z_dim: int, channels_img: int, features_g: int):
        super().__init__()
        f = features_g
        self.gen = nn.Sequential(
            _up(z_dim, 2 * f, 7, 1, 0),
            _up(2 * f, f, 4, 2, 1),
            nn.ConvTranspose2d(f, channels_img, 4, 2, 1),
            nn.Tanh(),
        )

    def forward(self, x):
        return self.gen(x)


class Discriminator(nn.Module):
    """[N,C,28,28] -> [N,1,1,1] sigmoid score  (src/utils/util_dcgan.py:103-125)."""

    def __init__(self, channels_img: int, features_d: int):
        super().__init__()
        f = features_d
        self.disc = nn.Sequential(
            nn.Conv2d(channels_img, f, 4, 2, 1),
            nn.LeakyReLU(0.2),
            _down(f, 2 * f, 4, 2, 1),
            nn.Conv2d(2 * f, 1, 7, 2, 0),
            nn.Sigmoid(),
        )

    def forward(self, x):
        return self.disc(x)


class Encoder(nn.Module):
    """Inverter encoder [N,C,28,28] -> [N,enc_dim,1,1]
    (src/inverter/utils_ae/util_inverter.py:164-184)."""

    def __init__(self, enc_dim: int, channels_img: int, features_e: int = 64):
        super().__init__()
        f = features_e
        self.enc = nn.Sequential(
            nn.Conv2d(channels_img, f, 4, 2, 1),
            nn.LeakyReLU(0.2),
            _down(f, 2 * f, 4, 2, 1),
            nn.Conv2d(2 * f, enc_dim, 7, 2, 0),
        )

    def forward(self, x):
        return self.enc(x)


class VQDecoderGAN(nn.Module):
    """`VectorQuantizedVAE_GAN.decoder` (src/inverter/utils_vq_vae/util_model.py:283-289):
    same stack as `Generator`, stored under `decoder.*`, xavier-uniform conv init with
    zero bias (`weights_init`, :39-46)."""

    def __init__(self, channels_img: int = 1, embedded_dim: int = 100, features_g: int = 64):
        super().__init__()
        f = features_g
        self.decoder = nn.Sequential(
            _up(embedded_dim, 2 * f, 7, 1, 0),
            _up(2 * f, f, 4, 2, 1),
            nn.ConvTranspose2d(f, channels_img, 4, 2, 1),
            nn.Tanh(),
        )
        for m in self.modules():
            if isinstance(m, (nn.ConvTranspose2d, nn.Conv2d)):
                nn.init.xavier_uniform_(m.weight.data)
                m.bias.data.fill_(0)

    def forward(self, x):
        return self.decoder(x)


class VQDecoderMNIST(nn.Module):
    """`VectorQuantizedVAE_MNIST.decoder` (src/inverter/utils_vq_vae/util_model.py:237-244): the BatchNorm-free variant,
    ConvT(D -> h, k7, s2, p0) ReLU ConvT(h -> h/2, k4, s2, p1) ReLU ConvT(h/2 -> C, k4, s2, p1) Tanh, stored under
    `decoder.{0,2,4}`; xavier-uniform weights and zero biases (`weights_init`, :39-46)."""

    def __init__(self, channels_img: int = 1, embedded_dim: int = 64, num_hiddens: int = 64):
        super().__init__()
        h = num_hiddens
        self.decoder = nn.Sequential(
            nn.ConvTranspose2d(embedded_dim, h, 7, 2, 0), nn.ReLU(True),
            nn.ConvTranspose2d(h, h // 2, 4, 2, 1), nn.ReLU(True),
            nn.ConvTranspose2d(h // 2, channels_img, 4, 2, 1), nn.Tanh())
        for m in self.modules():
            if isinstance(m, nn.ConvTranspose2d):
                nn.init.xavier_uniform_(m.weight.data)
                m.bias.data.fill_(0)

    def forward(self, x):
        return self.decoder(x)


class ResBlockBN(nn.Module):
    """`ResBlockBatchNorm` (src/inverter/utils_vq_vae/util_model.py:151-164): x + BN(Conv1x1(ReLU(BN(Conv3x3(ReLU(x)))))).
    The block's first ReLU is IN PLACE (`nn.ReLU(True)`), so by the time the sum is formed `x` already holds relu(x):
    the block computes relu(x) + block(relu(x)).  Restated explicitly here (no in-place op)."""

    def __init__(self, dim: int):
        super().__init__()
        self.block = nn.Sequential(nn.ReLU(), nn.Conv2d(dim, dim, 3, 1, 1), nn.BatchNorm2d(dim), nn.ReLU(),
                                   nn.Conv2d(dim, dim, 1), nn.BatchNorm2d(dim))

    def forward(self, x):
        r = torch.relu(x)
        return r + self.block[1:](r)


class VQGridDecoder(nn.Module):
    """`VectorQuantizedVAE.decoder` (src/inverter/utils_vq_vae/util_model.py:194-203): two ResBlockBatchNorm(dim) on the
    7x7 latent grid, ReLU, ConvT(dim, dim, 4, 2, 1) + BN + ReLU, ConvT(dim, C, 4, 2, 1) + Tanh; same state_dict keys
    (`decoder.{0,1}.block.{1,2,4,5}.*`, `decoder.{3,4,6}.*`); xavier-uniform conv weights, zero biases (:39-46)."""

    def __init__(self, channels_img: int = 1, embedded_dim: int = 64):
        super().__init__()
        d = embedded_dim
        self.decoder = nn.Sequential(ResBlockBN(d), ResBlockBN(d), nn.ReLU(), nn.ConvTranspose2d(d, d, 4, 2, 1), nn.BatchNorm2d(d),
                                     nn.ReLU(), nn.ConvTranspose2d(d, channels_img, 4, 2, 1), nn.Tanh())
        for m in self.modules():
            if isinstance(m, (nn.ConvTranspose2d, nn.Conv2d)):
                nn.init.xavier_uniform_(m.weight.data)
                m.bias.data.fill_(0)

    def forward(self, x):
        return self.decoder(x)


def pso_weights_codebook(data) -> torch.Tensor:
    """`pso_weights` (src/inverter/utils_vq_vae/util_model.py:49-54): the PSO particles' final positions (a float32 DataFrame,
    one row per particle, built by training/vq_vae.py:35-57) BECOME the codebook: nn.Embedding.weight = tensor(data.values)."""
    return torch.tensor(np.asarray(getattr(data, "values", data)))


def decode_snapped(decoder: nn.Module, x: torch.Tensor, codebook: torch.Tensor):
    """VQ-snapped variant of BASELINE config 3: positions -> nearest code (`vq`, util_function.py:4-23) -> codebook row ->
    `VectorQuantizedVAE_GAN.decoder`.  Returns (indices, images)."""
    idx = vq_nearest(x, codebook)
    with torch.no_grad():
        return idx, decoder(codebook[idx].view(x.shape[0], -1, 1, 1))


class G256(nn.Module):
    """EXTENSION: 1x256x256 generator built from the reference's `_block`
    (src/utils/util_dcgan.py:141-146) because the reference G is hard-wired to 28x28.
    ConvT(z->32f,k4,s1,p0)+BN+ReLU (4x4) -> five k4/s2/p1 blocks 32f->16f->8f->4f->2f->f
    -> ConvT(f->C,k4,s2,p1)+Tanh (256x256).  No reference counterpart: oracle == this."""

    def __init__(self, z_dim: int = 100, channels_img: int = 1, features_g: int = 64):
        super().__init__()
        f = features_g
        layers = [_up(z_dim, 32 * f, 4, 1, 0)]
        c = 32 * f
        for _ in range(5):
            layers.append(_up(c, c // 2, 4, 2, 1))
            c //= 2
        layers += [nn.ConvTranspose2d(c, channels_img, 4, 2, 1), nn.Tanh()]
        self.gen = nn.Sequential(*layers)

    def forward(self, x):
        return self.gen(x)


class Bottleneck(nn.Module):
    """`resnet_block` (src/pso/util_cnn.py:160-190): 1x1 (cin -> mid) BN ReLU, 3x3 stride s pad 1 BN ReLU,
    1x1 (mid -> 4 mid) BN, + shortcut (1x1 stride s conv + BN when the shape changes), ReLU.
    The stride sits on the 3x3 conv; no conv has a bias."""

    def __init__(self, cin: int, mid: int, stride: int, project: bool):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, mid, 1, 1, 0, bias=False)
        self.bn1 = nn.BatchNorm2d(mid)
        self.conv2 = nn.Conv2d(mid, mid, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(mid)
        self.conv3 = nn.Conv2d(mid, 4 * mid, 1, 1, 0, bias=False)
        self.bn3 = nn.BatchNorm2d(4 * mid)
        self.identity_downsample = nn.Sequential(nn.Conv2d(cin, 4 * mid, 1, stride, bias=False),
                                                 nn.BatchNorm2d(4 * mid)) if project else None

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = F.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        sc = x if self.identity_downsample is None else self.identity_downsample(x)
        return F.relu(y + sc)


class Resnet(nn.Module):
    """The reference's assessor (src/pso/util_cnn.py:81-133): 7x7 stride-2 pad-3 stem + BN + ReLU,
    3x3 stride-2 pad-1 max pool, four stages of bottlenecks (widths 64/128/256/512, the first block of
    stages 2-4 has stride 2), a global MAX pool (the attribute is named `avgpool`, :99) and
    `fc: 2048 -> n_class`.  `layers = [3, 4, 6, 3]` is ResNet-50.  Same state_dict keys as the reference."""

    def __init__(self, layers: Sequence[int], image_channels: int, iid_classes: Sequence[int], n_class: int = 2):
        super().__init__()
        self.conv1 = nn.Conv2d(image_channels, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.class_to_idx = {c: i for i, c in enumerate(sorted(iid_classes))}
        self.idx_to_class = {i: c for c, i in self.class_to_idx.items()}
        cin = 64
        for li, (n, mid, stride) in enumerate(zip(layers, (64, 128, 256, 512), (1, 2, 2, 2)), start=1):
            blocks = []
            for b in range(n):
                first = b == 0
                blocks.append(Bottleneck(cin, mid, stride if first else 1, first and (stride != 1 or cin != 4 * mid)))
                cin = 4 * mid
            setattr(self, f"layer{li}", nn.Sequential(*blocks))
        self.fc = nn.Linear(2048, n_class)

    def forward_avgpool(self, x):
        x = F.max_pool2d(F.relu(self.bn1(self.conv1(x))), 3, 2, 1)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return F.adaptive_max_pool2d(x, 1).flatten(1)

    def forward(self, x):
        return self.fc(self.forward_avgpool(x))


def seeded_fill(model: nn.Module, seed: int) -> None:
    """Deterministic 'trained-like' parameters, independent of module construction order: every
    state_dict entry is drawn in sorted-key order from one seeded generator.  Conv weights
    N(0, 2 / fan_in), BN gamma U(0.7, 1.3) (the last BN of a bottleneck U(0.1, 0.4): keeps the residual
    stream O(1) over 16 blocks), beta / running_mean N(0, 0.1), running_var U(0.5, 1.5), fc so that
    the logits are O(1).  Used on the reference model by oracle/gen_golden_resnet.py and on the port / the
    CUDA path by the tests, so that a 94 MB state_dict need not be committed."""
    gen = torch.Generator().manual_seed(seed)
    sd = model.state_dict()
    with torch.no_grad():
        for k in sorted(sd.keys()):
            v = sd[k]
            if k.endswith("num_batches_tracked"):
                continue
            if v.dim() == 4:
                fan_in = v.shape[1] * v.shape[2] * v.shape[3]
                v.copy_(torch.randn(v.shape, generator=gen) * math.sqrt(2.0 / fan_in))
            elif k.endswith("running_var"):
                v.copy_(torch.rand(v.shape, generator=gen) + 0.5)
            elif k.endswith("running_mean"):
                v.copy_(torch.randn(v.shape, generator=gen) * 0.1)
            elif k.startswith("fc."):
                v.copy_(torch.randn(v.shape, generator=gen) * (0.07 if v.dim() == 2 else 0.1))
            elif k.endswith(".weight"):       # BN gamma
                lo, hi = (0.1, 0.4) if ".bn3." in k else (0.7, 1.3)
                v.copy_(torch.rand(v.shape, generator=gen) * (hi - lo) + lo)
            else:                             # BN beta
                v.copy_(torch.randn(v.shape, generator=gen) * 0.1)


def initialize_weights(model: nn.Module) -> None:
    """N(0, 0.02) on conv, conv-transpose *and BN gamma* (src/utils/util_dcgan.py:45-48)."""
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d, nn.BatchNorm2d)):
            nn.init.normal_(m.weight.data, 0.0, 0.02)


def randomize_bn_stats(model: nn.Module, gen: torch.Generator) -> None:
    """'Trained-like' BN statistics so that BN-folding bugs are visible."""
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d):
            n = m.num_features
            m.running_mean.copy_(torch.randn(n, generator=gen) * 0.3)
            m.running_var.copy_(torch.rand(n, generator=gen) * 1.5 + 0.25)
            m.weight.data.copy_(torch.rand(n, generator=gen) * 1.0 + 0.5)
            m.bias.data.copy_(torch.randn(n, generator=gen) * 0.2)


# --------------------------------------------------------------------------------------
# Fitness stage
# --------------------------------------------------------------------------------------


def rescale_minmax(img: torch.Tensor) -> torch.Tensor:
    """(img - min) / (max - min), min/max over the WHOLE tensor
    (src/utils/util_data.py:129-135).  Per-image only because the reference calls it with
    batch 1; use `rescale_minmax_per_image` for a batch."""
    lo, hi = torch.min(img), torch.max(img)
    return (img - lo) / (hi - lo)


def rescale_minmax_per_image(img: torch.Tensor) -> torch.Tensor:
    flat = img.reshape(img.shape[0], -1)
    lo = flat.min(dim=1).values.view(-1, *([1] * (img.dim() - 1)))
    hi = flat.max(dim=1).values.view(-1, *([1] * (img.dim() - 1)))
    return (img - lo) / (hi - lo)


def score_to_fitness(p: torch.Tensor, mode: str, thr: float = 0.0, eps: float = 0.1) -> torch.Tensor:
    """min(p+thr,1)+eps  |  1-min(p+thr,1)+eps   (src/pso/util_discovery.py:75-80)."""
    clipped = torch.minimum(p + thr, torch.ones((), dtype=p.dtype))
    if mode == "optimize_in_training":
        return clipped + eps
    if mode == "optimize_out_training":
        return (1.0 - clipped) + eps
    raise ValueError(mode)


@torch.no_grad()
def particles_to_img(model_gan: nn.Module, pos: torch.Tensor, dim_space: int):
    """One particle -> (img_rescaled, img)   (src/pso/util_discovery.py:33-50)."""
    z = pos.reshape(1, dim_space, 1, 1).to(torch.float32)
    model_gan.eval()
    img = model_gan(z)
    return rescale_minmax(img), img


@torch.no_grad()
def fitness_assessor(model_gan, model_cnn, pos, dim_space, *, mode="optimize_out_training",
                     class_index: Optional[int] = None, thr=0.0, eps=0.1):
    """`Discovery.fitness` for ONE particle (src/pso/util_discovery.py:52-82): G, min-max
    rescale, assessor logits, softmax, class pick (index if n_class>2 else column 1),
    clamp/affine.  Returns (np.float32[1], img_rescaled, img)."""
    if not isinstance(pos, torch.Tensor):
        pos = torch.from_numpy(np.asarray(pos))
    img_rescaled, img = particles_to_img(model_gan, pos, dim_space)
    assert torch.max(img_rescaled) <= 1.0 and torch.min(img_rescaled) >= 0.0
    model_cnn.eval()
    out = model_cnn(img_rescaled.float())
    prob = F.softmax(out, dim=1, dtype=torch.float32)[0]
    p = prob[class_index] if prob.shape[0] > 2 else prob[1]
    fit = score_to_fitness(p.reshape(1), mode, thr, eps)
    return fit.numpy().astype(np.float32), img_rescaled, img


@torch.no_grad()
def fitness_dscore(model_gan, model_disc, pos, dim_space, *, mode="optimize_out_training",
                   thr=0.0, eps=0.1):
    """EXTENSION "discriminator-score fitness" (SURVEY.md §8 a17): p = D(G(z)) on the raw
    [-1,1] image (the range D was trained on, src/utils/util_dcgan.py:193-197), then the
    clamp/affine of src/pso/util_discovery.py:75-80.  Works for one particle or a batch
    (there is no whole-tensor rescale on this path).  Returns (fp32[P], img)."""
    if not isinstance(pos, torch.Tensor):
        pos = torch.from_numpy(np.asarray(pos))
    z = pos.reshape(-1, dim_space, 1, 1).to(torch.float32)
    model_gan.eval()
    model_disc.eval()
    img = model_gan(z)
    p = model_disc(img).reshape(-1)
    return score_to_fitness(p, mode, thr, eps).numpy().astype(np.float32), img


@torch.no_grad()
def fitness_inverter(model_gan, pos, target, dim_space, *, assessor_fitness=None,
                     eps=0.1, w_ass=1.0, w_rec=1.0):
    """`Discovery.fitness_pso_inverter` (src/pso/util_discovery.py:84-101):
    w_ass * fitness(pos) + w_rec * mean((target - G(pos))**2) + eps, where fitness(pos)
    already contains one eps (eps is added twice) and the MSE uses the RAW tanh image.
    `assessor_fitness` is the value of `fitness(pos)`; pass 0 (with w_ass=0) for the pure
    reconstruction objective of BASELINE config 3.  Batched over particles."""
    if not isinstance(pos, torch.Tensor):
        pos = torch.from_numpy(np.asarray(pos))
    z = pos.reshape(-1, dim_space, 1, 1).to(torch.float32)
    model_gan.eval()
    img = model_gan(z)
    mse = ((target.float() - img.float()) ** 2).reshape(img.shape[0], -1).mean(dim=1)
    ass = 0.0 if assessor_fitness is None else torch.as_tensor(assessor_fitness, dtype=torch.float32)
    fit = w_ass * ass + w_rec * mse + eps
    return fit.numpy().astype(np.float32), img


# --------------------------------------------------------------------------------------
# PSO arithmetic
# --------------------------------------------------------------------------------------


def pso_move(pos, vel, pbest_pos, gbest_pos, r1, r2, w, w_cogn, w_soci,
             v_max=math.inf, lo=-math.inf, hi=math.inf, dtype=np.float64):
    """One `move_particles` over the whole swarm, vectorised, in the reference's operation
    order (src/pso/util_pso.py:39-50):
        inertia   = w * vel
        best_cogn = w_cogn * r1 * (gbest - pos)      # NB: "cogn" pulls to the GLOBAL best
        best_soci = w_soci * r2 * (pbest - pos)      #     "soci" pulls to the PERSONAL best
        vel = (inertia + best_soci) + best_cogn ;  pos = pos + vel
    r1, r2 are ONE scalar per particle.  Extensions (default off): velocity clamp to
    [-v_max, v_max] after the update, then position clamp to [lo, hi].
    dtype=float64 is the reference; dtype=float32 is the same-order fp32 restatement."""
    pos = np.asarray(pos, dtype=dtype)
    vel = np.asarray(vel, dtype=dtype)
    pb = np.asarray(pbest_pos, dtype=dtype)
    gb = np.asarray(gbest_pos, dtype=dtype)[None, :]
    r1 = np.asarray(r1, dtype=dtype)[:, None]
    r2 = np.asarray(r2, dtype=dtype)[:, None]
    w, w_cogn, w_soci = dtype(w), dtype(w_cogn), dtype(w_soci)
    inertia = w * vel
    best_cogn = (w_cogn * r1) * (gb - pos)
    best_soci = (w_soci * r2) * (pb - pos)
    new_vel = (inertia + best_soci) + best_cogn
    if math.isfinite(v_max):
        new_vel = np.clip(new_vel, dtype(-v_max), dtype(v_max))
    new_pos = pos + new_vel
    if math.isfinite(lo) or math.isfinite(hi):
        new_pos = np.clip(new_pos, dtype(lo), dtype(hi))
    return new_pos, new_vel


def pbest_update(fit, pbest_val):
    """Strict `<` personal-best rule (src/pso/util_pso.py:122-124).  NaN never improves.
    Returns (new_pbest_val, improved_mask)."""
    fit = np.asarray(fit, dtype=np.float32)
    pbest_val = np.asarray(pbest_val, dtype=np.float32)
    improved = fit < pbest_val
    return np.where(improved, fit, pbest_val).astype(np.float32), improved


def gbest_scan(pbest_val, incumbent_val):
    """`update_global_best` scan (src/pso/util_pso.py:135-151): strict `<` in index order
    starting from the incumbent, so the lowest index wins ties and a value equal to the
    incumbent never replaces it.  Returns (value, index) with index -1 when unchanged."""
    best, idx = np.float32(incumbent_val), -1
    for i, v in enumerate(np.asarray(pbest_val, dtype=np.float32)):
        if v < best:
            best, idx = v, i
    return best, idx


def f32_order_key(x) -> np.ndarray:
    """Order-preserving float32 -> uint32 map (sign-flip trick) used by the packed
    (value,index) min-loc key; NaN sorts above +inf; -0.0 and +0.0 map to the same key, like the
    reference's `<` treats them (src/pso/util_pso.py:139-143).  """
    b = np.asarray(x, dtype=np.float32).view(np.uint32).copy()
    b[...] = np.where(b == np.uint32(0x80000000), np.uint32(0), b)
    neg = (b >> 31).astype(bool)
    return np.where(neg, ~b, b | np.uint32(0x80000000)).astype(np.uint32)


def pack_minloc_key(val, index) -> np.ndarray:
    return (f32_order_key(val).astype(np.uint64) << np.uint64(32)) | np.asarray(index, dtype=np.uint64)


def mean_pairwise_distance(pos) -> float:
    """`Swarm.mse` (src/pso/util_pso.py:76-86): mean Euclidean distance over all
    unordered pairs (the name notwithstanding)."""
    pos = np.asarray(pos, dtype=np.float64)
    n = pos.shape[0]
    if n < 2:
        return float("nan")
    g = pos @ pos.T
    sq = np.diag(g)
    d2 = np.maximum(sq[:, None] + sq[None, :] - 2 * g, 0.0)
    iu = np.triu_indices(n, 1)
    # recompute exactly for the returned pairs to avoid cancellation error
    d = np.linalg.norm(pos[iu[0]] - pos[iu[1]], axis=1) if n <= 2048 else np.sqrt(d2[iu])
    return float(d.mean())


# --------------------------------------------------------------------------------------
# Philox4x32-10 (EXTENSION: in-kernel counter-based draws; SURVEY.md §8 "Philox draws")
# --------------------------------------------------------------------------------------

_PH_M0, _PH_M1 = 0xD2511F53, 0xCD9E8D57
_PH_W0, _PH_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10 (Salmon et al. 2011).  All arguments uint32 arrays/scalars."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) for c in (c0, c1, c2, c3))
    k0 = np.uint64(k0)
    k1 = np.uint64(k1)
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(_PH_M0) * c0
        p1 = np.uint64(_PH_M1) * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & mask
        hi1, lo1 = p1 >> np.uint64(32), p1 & mask
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & mask, lo1, (hi0 ^ c3 ^ k1) & mask, lo0
        k0 = (k0 + np.uint64(_PH_W0)) & mask
        k1 = (k1 + np.uint64(_PH_W1)) & mask
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def philox_r1r2(seed: int, iteration: int, first_index: int, count: int):
    """Per-particle (r1, r2) in [0,1): counter = (global particle index, iteration, 0, 0),
    key = (seed_lo, seed_hi); r1 = (x0>>8)*2^-24, r2 = (x1>>8)*2^-24."""
    idx = np.arange(first_index, first_index + count, dtype=np.uint64) & np.uint64(0xFFFFFFFF)
    x0, x1, _, _ = philox4x32_10(idx, np.full(count, iteration, dtype=np.uint64), 0 * idx, 0 * idx,
                                 seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    scale = np.float32(2.0 ** -24)
    return (x0 >> np.uint32(8)).astype(np.float32) * scale, (x1 >> np.uint32(8)).astype(np.float32) * scale


# --------------------------------------------------------------------------------------
# VQ nearest-codebook ("next" row; src/inverter/utils_vq_vae/util_function.py:4-23)
# --------------------------------------------------------------------------------------


@torch.no_grad()
def vq_nearest(x: torch.Tensor, codebook: torch.Tensor) -> torch.Tensor:
    c2 = torch.sum(codebook ** 2, dim=1)
    x2 = torch.sum(x ** 2, dim=1, keepdim=True)
    d = torch.addmm(c2 + x2, x, codebook.t(), alpha=-2.0, beta=1.0)
    return torch.min(d, dim=1).indices


# --------------------------------------------------------------------------------------
# The faithful per-particle loop (cpu_baseline "port"): same structure, dtypes and
# operation order as src/pso/util_pso.py — python loop over particles, numpy fp64 state,
# batch-1 objective, global numpy RNG drawn as r1 then r2 per particle.
# --------------------------------------------------------------------------------------


@dataclass
class PortParticle:
    pos: np.ndarray
    vel: np.ndarray
    p_best_pos: np.ndarray
    p_best_val: float = np.inf
    img: Optional[torch.Tensor] = None


@dataclass
class PortSwarm:
    """objective(particle) -> float32 scalar/array; mirrors Swarm.optimize
    (src/pso/util_pso.py:167-193) minus plotting / tensorboard I/O."""
    objective: Callable[[PortParticle], np.ndarray]
    num_particles: int
    n_iterations: int
    dim_space: int
    tolerance: float = 1e-4
    w_inertia: float = 0.5
    w_cogn: float = 0.8
    w_soci: float = 0.9
    compute_diversity: bool = True
    swarm: List[PortParticle] = field(default_factory=list)

    def __post_init__(self):
        self.g_best_pos = np.empty(self.dim_space)
        self.g_best_val = [np.inf]
        self.g_best_val_dummy = []

    def init_particles(self, positions: Optional[Sequence[np.ndarray]] = None, images=None):
        # src/pso/util_pso.py:25-37,88-91 — pos first, then vel, particle by particle
        for i in range(self.num_particles):
            if positions is not None:
                pos = positions[i]
            else:
                pos = np.random.randn(self.dim_space)
            vel = (np.random.randn(self.dim_space) - 0.5) / 10
            self.swarm.append(PortParticle(pos=pos, vel=vel, p_best_pos=pos,
                                           img=None if images is None else images[i]))
        return np.asarray([p.pos for p in self.swarm])

    def update_personal_best(self):
        for p in self.swarm:
            f = self.objective(p)
            if f < p.p_best_val:
                p.p_best_val = f
                p.p_best_pos = p.pos

    def update_global_best(self):
        cand_val, cand_pos = self.g_best_val[-1], self.g_best_pos
        for p in self.swarm:
            if p.p_best_val < cand_val:
                cand_val, cand_pos = p.p_best_val, p.p_best_pos
        if cand_val < self.g_best_val[-1]:
            if self.g_best_val[-1] == np.inf:
                self.g_best_val[-1] = cand_val
            else:
                self.g_best_val.append(cand_val)
            self.g_best_pos = cand_pos
        self.g_best_val_dummy.append(cand_val)

    def move_particles(self):
        for p in self.swarm:
            inertia = self.w_inertia * p.vel
            r1 = np.random.random()
            best_cogn = self.w_cogn * r1 * (self.g_best_pos - p.pos)
            r2 = np.random.random()
            best_soci = self.w_soci * r2 * (p.p_best_pos - p.pos)
            p.vel = inertia + best_soci + best_cogn
            p.pos = p.pos + p.vel

    def diversity(self):
        # src/pso/util_pso.py:76-86 (python O(P^2) loop in the reference)
        d = []
        for i, a in enumerate(self.swarm):
            for b in self.swarm[i + 1:]:
                d.append(np.linalg.norm(a.pos - b.pos))
        return np.mean(d)

    def optimize(self, schedule_inertia=False, early_stopping=True, positions=None, images=None):
        self.init_particles(positions, images)
        i = 1
        history = {"mean_mse": [], "global_best_val": []}
        while i < self.n_iterations + 1:
            self.update_personal_best()
            self.update_global_best()
            if i > 1 and schedule_inertia:
                self.w_inertia = 0.99 * self.w_inertia
            self.move_particles()
            if self.compute_diversity:
                history["mean_mse"].append(self.diversity())
            if i > 2 and len(self.g_best_val) > 2 and early_stopping:
                if abs(self.g_best_val[-1] - self.g_best_val[-2]) < self.tolerance:
                    break
            i += 1
        return history, i


def make_assessor_objective(model_gan, model_cnn, dim_space, mode="optimize_out_training",
                            class_index=None, thr=0.0, eps=0.1):
    def obj(p: PortParticle):
        return fitness_assessor(model_gan, model_cnn, p.pos, dim_space, mode=mode,
                                class_index=class_index, thr=thr, eps=eps)[0]
    return obj


def make_dscore_objective(model_gan, model_disc, dim_space, mode="optimize_out_training", thr=0.0, eps=0.1):
    def obj(p: PortParticle):
        return fitness_dscore(model_gan, model_disc, p.pos, dim_space, mode=mode, thr=thr, eps=eps)[0]
    return obj


def make_recon_objective(model_gan, dim_space, eps=0.1, w_rec=1.0):
    def obj(p: PortParticle):
        return fitness_inverter(model_gan, p.pos, p.img, dim_space, eps=eps, w_ass=0.0, w_rec=w_rec)[0]
    return obj


# --------------------------------------------------------------------------------------
# Batched fp32 torch-CPU RESTATEMENT of one PSO iteration with the discriminator-score fitness
# (BASELINE.md section 4 step 4).  The reference itself cannot run this workload (a Discriminator as
# `model_cnn` raises IndexError at src/pso/util_discovery.py:74) and evaluates one particle at a
# time; this is the like-for-like CPU figure for BASELINE configs 2 / 5: the SAME arithmetic as
# `PortSwarm` with `make_dscore_objective`, vectorised over the swarm -- batched G + D forward
# (`fitness_dscore` on chunks), vectorised strict-< personal best, lowest-index argmin, vectorised
# move with the Philox draws of SURVEY.md section 8.  Always labelled "restatement", never "reference".
# --------------------------------------------------------------------------------------

class BatchedDScoreSwarm:
    def __init__(self, model_gan, model_disc, num_particles, dim_space, w_inertia=0.5, w_cogn=0.8, w_soci=0.9,
                 seed=42, chunk=4096, mode="optimize_out_training"):
        self.gen, self.disc, self.P, self.D = model_gan.eval(), model_disc.eval(), num_particles, dim_space
        self.w, self.c_cogn, self.c_soci, self.seed, self.chunk, self.mode = w_inertia, w_cogn, w_soci, seed, chunk, mode
        rng = np.random.RandomState(seed)
        self.pos = rng.randn(num_particles, dim_space).astype(np.float32)
        self.vel = ((rng.randn(num_particles, dim_space) - 0.5) / 10).astype(np.float32)
        self.pbest_pos = self.pos.copy()
        self.pbest_val = np.full(num_particles, np.inf, np.float32)
        self.gbest_val, self.gbest_pos = np.float32(np.inf), np.zeros(dim_space, np.float32)
        self.iteration = 0

    @torch.no_grad()
    def step(self):
        fit = np.empty(self.P, np.float32)
        for a in range(0, self.P, self.chunk):
            b = min(self.P, a + self.chunk)
            fit[a:b] = fitness_dscore(self.gen, self.disc, torch.from_numpy(self.pos[a:b]), self.D, mode=self.mode)[0]
        self.pbest_val, improved = pbest_update(fit, self.pbest_val)
        self.pbest_pos[improved] = self.pos[improved]
        j = int(np.argmin(self.pbest_val))                     # first occurrence == the reference's in-order scan
        if self.pbest_val[j] < self.gbest_val:
            self.gbest_val, self.gbest_pos = self.pbest_val[j], self.pbest_pos[j].copy()
        self.iteration += 1
        r1, r2 = philox_r1r2(self.seed, self.iteration, 0, self.P)
        self.pos, self.vel = pso_move(self.pos, self.vel, self.pbest_pos, self.gbest_pos, r1, r2, self.w, self.c_cogn,
                                      self.c_soci, dtype=np.float32)
        return fit
