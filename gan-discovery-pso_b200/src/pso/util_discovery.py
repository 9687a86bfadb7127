"""`Discovery` — objective layer of the PSO loop with the reference's interface
(src/pso/util_discovery.py:12-106), evaluated for the WHOLE swarm at once on the B200 engine.

Single-particle calls (`fitness(dim_space, pos)`) keep the reference's return shapes; the
batched forms (`*_batch`) are what `Swarm` uses.  Differences that are deliberate:
  * min-max rescale is per image (the reference's whole-tensor `rescale_torch`,
    src/utils/util_data.py:129-135, is per image only because its batch is 1);
  * the two range asserts of :63-64 (host syncs per particle) are dropped;
  * a `src.pso.util_cnn.Resnet` assessor (the reference's own, :52-82) runs on the engine; any other
    `model_cnn` is called as a torch module on the rescaled batch;
  * passing a `Discriminator` as `model_cnn` selects the discriminator-score fitness
    (extension; in the reference it raises IndexError at :74).
"""
from __future__ import annotations

import numpy as np
import torch

from psob import engine as _engine
from psob import _ffi
from src.utils import util_dcgan
from src.pso import util_cnn

_MODES = ("optimize_in_training", "optimize_out_training")


class Discovery:
    def __init__(self, iid_class, model_gan, model_cnn, device, control_pso_fitness="optimize_out_training",
                 iid_classes=None, obj_fun_threshold=0.0):
        self.model_gan = model_gan
        self.model_cnn = model_cnn
        self.device = torch.device(device)
        self.threshold = obj_fun_threshold
        self.iid_class = iid_class
        self.control_pso_fitness = control_pso_fitness
        if iid_classes is not None:
            self.class_to_idx = {c: i for i, c in enumerate(sorted(iid_classes))}
            self.idx_to_class = {i: c for c, i in self.class_to_idx.items()}
        self._engines = {}

    # -- engine plumbing -----------------------------------------------------------------------
    def _class_index(self):
        if hasattr(self, "class_to_idx") and self.iid_class in self.class_to_idx:
            return self.class_to_idx[self.iid_class]
        return None

    def generator_net(self, P) -> _engine.GeneratorNet:
        if not isinstance(self.model_gan, util_dcgan.Generator):
            raise _ffi.PsobError("model_gan must be src.utils.util_dcgan.Generator (B200 engine); load the reference "
                                 "checkpoint into it with load_state_dict")
        return self.model_gan.engine_net(P)

    def assessor_fitness(self, P, eps=0.1):
        """The fitness stage for `fitness()`: D-score (extension) or assessor CNN + softmax."""
        if self.control_pso_fitness not in _MODES:
            raise ValueError(self.control_pso_fitness)
        if isinstance(self.model_cnn, util_dcgan.Discriminator):
            return _engine.DScoreFitness(self.model_cnn.engine_net(P), self.control_pso_fitness, self.threshold, eps)
        if isinstance(self.model_cnn, util_cnn.Resnet):       # the reference's assessor, on the engine
            return _engine.ResnetFitness(self.model_cnn.engine_net(P), self._class_index(), self.control_pso_fitness, self.threshold, eps)
        self.model_cnn.eval()
        return _engine.AssessorFitness(self.model_cnn, self._class_index(), self.control_pso_fitness, self.threshold, eps)

    # -- batched API ---------------------------------------------------------------------------
    @torch.no_grad()
    def particles_to_img_batch(self, pos: torch.Tensor):
        """pos [P, D] (cuda) -> (img_rescaled [P,1,28,28], img [P,1,28,28])."""
        P = pos.shape[0]
        net = self.generator_net(P)
        z = torch.zeros(P, net.in_pitch, dtype=torch.bfloat16, device=pos.device)
        _engine.pack_latent(pos.reshape(P, -1).float().contiguous(), z)
        net.forward(z)
        img = net.image()
        resc = torch.empty_like(img)
        _engine.rescale_minmax(img, resc)
        return resc.view(P, 1, 28, 28), img.view(P, 1, 28, 28).clone()

    @torch.no_grad()
    def fitness_batch(self, pos: torch.Tensor, eps=0.1):
        """-> (fitness fp32 [P] on device, img_rescaled, img)."""
        P = pos.shape[0]
        img_rescaled, img = self.particles_to_img_batch(pos)
        fit = torch.empty(P, dtype=torch.float32, device=pos.device)
        self.assessor_fitness(P, eps)(self.generator_net(P), fit)
        return fit, img_rescaled, img

    @torch.no_grad()
    def fitness_pso_inverter_batch(self, pos, img, eps=0.1, w_ass=1.0, w_rec=1.0):
        P = pos.shape[0]
        fit_ass, img_rescaled, img_rec = self.fitness_batch(pos, eps)
        fit = torch.empty(P, dtype=torch.float32, device=pos.device)
        tgt = img.to(pos.device).float().reshape(P, -1).contiguous()
        _engine.recon_fitness(self.generator_net(P).image(), tgt, fit, fit_ass, w_ass, w_rec, eps)
        return fit, img_rescaled, img_rec

    # -- reference-shaped single-particle API ---------------------------------------------------
    def _as_row(self, dim_space, pos):
        if not isinstance(pos, torch.Tensor):
            pos = torch.from_numpy(np.asarray(pos))
        return pos.to(self.device).reshape(1, dim_space).to(torch.float32)

    def particles_to_img(self, dim_space, pos):
        return self.particles_to_img_batch(self._as_row(dim_space, pos))

    def fitness(self, dim_space, pos, eps=0.1):
        fit, img_rescaled, img = self.fitness_batch(self._as_row(dim_space, pos), eps)
        return fit.detach().cpu().numpy(), img_rescaled, img

    def fitness_pso_inverter(self, dim_space, pos, img, eps=0.1, w_ass=1.0, w_rec=1.0):
        if not isinstance(img, torch.Tensor):
            img = torch.from_numpy(np.asarray(img))
        fit, img_rescaled, img_rec = self.fitness_pso_inverter_batch(self._as_row(dim_space, pos), img.reshape(1, -1), eps, w_ass, w_rec)
        return fit.detach().cpu().numpy(), img_rescaled, img_rec

    def load_gan(self):
        pass

    def load_cnn(self):
        pass
