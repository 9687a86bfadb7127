"""`Resnet` / `resnet_block` / `get_cnn` with the reference's constructor signatures and state_dict keys
(src/pso/util_cnn.py:23-38,81-190), whose forward runs on the B200 pixel-GEMM engine (`psob.engine.AssessorNet`).

The modules own ordinary fp32 parameters, so the reference's assessor checkpoint loads unchanged
(`model.load_state_dict(torch.load(...))`, src/training/pso_discovery.py:177-186).  Eval semantics only (BatchNorm
folded), single-channel 28x28 inputs (the PSO configs); `forward` returns logits fp32 [N, n_class].  Training code
(`train_model`, optimisers, AlexNet) is outside the swarm-evaluation scope.  There is no CPU path."""
from __future__ import annotations

import torch
import torch.nn as nn

from psob import engine as _engine
from psob import _ffi


def get_initializer(initializer_name, w):
    if initializer_name == "glorot_normal":
        return nn.init.xavier_normal_(tensor=w, gain=1.0)
    if initializer_name == "he_normal":
        return nn.init.kaiming_normal_(tensor=w, a=0, mode="fan_in", nonlinearity="leaky_relu")
    if initializer_name == "random_normal":
        return nn.init.normal_(tensor=w, mean=0.0, std=0.02)
    raise ValueError(initializer_name)


def initialize_weights(model, initializer_name):
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            get_initializer(initializer_name, m.weight.data)


class resnet_block(nn.Module):
    """Parameter holder of one bottleneck (src/pso/util_cnn.py:160-190); executed as part of `Resnet.forward`."""

    def __init__(self, in_channels, intermediate_channels, identity_downsample=None, stride=1):
        super().__init__()
        self.expansion = 4
        self.conv1 = nn.Conv2d(in_channels, intermediate_channels, kernel_size=1, stride=1, padding=0, bias=False)
        self.bn1 = nn.BatchNorm2d(intermediate_channels)
        self.conv2 = nn.Conv2d(intermediate_channels, intermediate_channels, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(intermediate_channels)
        self.conv3 = nn.Conv2d(intermediate_channels, intermediate_channels * self.expansion, kernel_size=1, stride=1, padding=0, bias=False)
        self.bn3 = nn.BatchNorm2d(intermediate_channels * self.expansion)
        self.identity_downsample = identity_downsample


class Resnet(nn.Module):
    def __init__(self, resnet_block, layers, image_channels, iid_classes, n_class=2):
        super().__init__()
        if image_channels != 1:
            raise _ffi.PsobError("the B200 assessor path implements single-channel images (the reference's PSO configs)")
        self.in_channels = 64
        self.conv1 = nn.Conv2d(image_channels, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.class_to_idx = {c: i for i, c in enumerate(sorted(iid_classes))}
        self.idx_to_class = {i: c for c, i in self.class_to_idx.items()}
        self.layer1 = self._make_layer(resnet_block, layers[0], intermediate_channels=64, stride=1)
        self.layer2 = self._make_layer(resnet_block, layers[1], intermediate_channels=128, stride=2)
        self.layer3 = self._make_layer(resnet_block, layers[2], intermediate_channels=256, stride=2)
        self.layer4 = self._make_layer(resnet_block, layers[3], intermediate_channels=512, stride=2)
        self.fc = nn.Linear(512 * 4, n_class)

    def _make_layer(self, resnet_block, num_residual_blocks, intermediate_channels, stride):
        shortcut = None
        if stride != 1 or self.in_channels != intermediate_channels * 4:
            shortcut = nn.Sequential(nn.Conv2d(self.in_channels, intermediate_channels * 4, kernel_size=1, stride=stride, bias=False),
                                     nn.BatchNorm2d(intermediate_channels * 4))
        blocks = [resnet_block(self.in_channels, intermediate_channels, shortcut, stride)]
        self.in_channels = intermediate_channels * 4
        blocks += [resnet_block(self.in_channels, intermediate_channels) for _ in range(num_residual_blocks - 1)]
        return nn.Sequential(*blocks)

    # -- engine plumbing ---------------------------------------------------------------------------
    def _version(self):
        return tuple(int(t._version) for t in self.state_dict().values())

    def engine_net(self, batch: int) -> _engine.AssessorNet:
        dev = next(self.parameters()).device
        key = (batch, self._version(), str(dev))
        cache = self.__dict__.setdefault("_nets", {})
        if key not in cache:
            cache.clear()
            cache[key] = _engine.AssessorNet(self.state_dict(), batch, device=dev)
        return cache[key]

    @torch.no_grad()
    def forward(self, x):
        """x: [N, 1, 28, 28] fp32 on the CUDA device (the min-max rescaled image) -> logits fp32 [N, n_class]."""
        if not x.is_cuda:
            raise _ffi.PsobError("Resnet.forward needs CUDA tensors (B200); there is no CPU fallback")
        if tuple(x.shape[1:]) != (1, 28, 28):
            raise _ffi.PsobError(f"Resnet.forward: expected [N, 1, 28, 28], got {tuple(x.shape)}")
        n = x.shape[0]
        return self.engine_net(n).forward(x.reshape(n, 784).float().contiguous()).clone()


def get_cnn(model_name, image_channels, iid_classes, n_class, img_dim=64, cnn_args=None):
    depth = {"ResNet50": [3, 4, 6, 3], "ResNet101": [3, 4, 23, 3], "ResNet152": [3, 8, 36, 3]}
    if model_name in depth:
        return Resnet(resnet_block, depth[model_name], image_channels, iid_classes, n_class)
    raise ValueError(model_name)
