"""`Particle` / `Swarm` with the reference's interface (src/pso/util_pso.py:24-193), running the
whole swarm as SoA tensors on the B200 engine (psob.engine.SwarmEngine).

Kept: constructor arguments, attribute names (`g_best_pos`, `g_best_val`, `g_best_val_dummy`,
`swarm`, `w_inertia`), the `optimize()` return tuple, strict-< best rules, the swapped naming
(`w_cogn` pulls to the global best, `w_soci` to the personal best, :46,48), the inertia
schedule and the early-stopping rule on the last two *improvements* (:186-188).
Changed on purpose: r1/r2 come from in-kernel Philox (or injected tensors) instead of numpy's
global MT19937; per-iteration PNG / TensorBoard image dumps are opt-in (`log_images_every`);
state is fp32 on device, exported as fp64 numpy on access; history is opt-in above 1 GiB.
"""
from __future__ import annotations

import math
import os

import numpy as np
import pandas as pd
import torch

from psob import engine as _engine
from psob import _ffi


class Particle:
    """Stand-alone (numpy, like the reference) or a view of row `index` of a Swarm's device state."""

    def __init__(self, dim, pos=None, img=None, _owner=None, _index=None):
        self._owner, self._index = _owner, _index
        if _owner is not None:
            self.img = img
            return
        if pos is not None:
            self.pos = pos
            self.img = img
        else:
            self.pos = np.random.randn(dim)
        self.vel = (np.random.randn(dim) - 0.5) / 10
        self.p_best_pos = self.pos
        self.p_best_val = np.inf
        self.history = [self.pos]
        self.history_vel = [self.vel]

    def __getattr__(self, name):
        owner = self.__dict__.get("_owner")
        if owner is None:
            raise AttributeError(name)
        i = self.__dict__["_index"]
        if name == "pos":
            return owner.engine.pos[i, :owner.dim_space].double().cpu().numpy()
        if name == "vel":
            return owner.engine.vel[i, :owner.dim_space].double().cpu().numpy()
        if name == "p_best_pos":
            owner.engine.flush_pbest()        # the improved position is copied by the next move; fold it in before looking
            return owner.engine.pbest_pos[i, :owner.dim_space].double().cpu().numpy()
        if name == "p_best_val":
            return owner.engine.pbest_val[i:i + 1].cpu().numpy()
        if name == "history":
            return [h for h in owner._history_of(i, 0)]
        if name == "history_vel":
            return [h for h in owner._history_of(i, 1)]
        raise AttributeError(name)

    def update_position(self):
        if self._owner is not None:
            raise _ffi.PsobError("device-backed particles move together: call Swarm.move_particles()")
        self.pos = self.pos + self.vel
        self.history.append(self.pos)

    def update_velocity(self, w_inertia, w_cogn, w_soci, g_best_pos):
        if self._owner is not None:
            raise _ffi.PsobError("device-backed particles move together: call Swarm.move_particles()")
        inertia = w_inertia * self.vel
        r1 = np.random.random()
        best_cogn = w_cogn * r1 * (g_best_pos - self.pos)
        r2 = np.random.random()
        best_soci = w_soci * r2 * (self.p_best_pos - self.pos)
        self.vel = inertia + best_soci + best_cogn
        self.history_vel.append(self.vel)


class _ParticleViews:
    """`Swarm.swarm`: a read-only sequence of `Particle` views onto the engine's SoA state, created on access (a list of
    a million Python objects would cost more than the swarm's first iterations)."""

    def __init__(self, owner, n):
        self._owner, self._n, self._imgs = owner, n, {}

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(self._n))]
        if i < 0:
            i += self._n
        if not 0 <= i < self._n:
            raise IndexError(i)
        return Particle(self._owner.dim_space, img=self._imgs.get(i), _owner=self._owner, _index=i)

    def __iter__(self):
        return (self[i] for i in range(self._n))


class Swarm:
    def __init__(self, plot_training_dir, discovery, num_particles, n_iterations, dim_space, device, tolerance=0.0001,
                 w_inertia=0.5, w_cogn=0.8, w_soci=0.9, *, seed=42, record_history=None, v_max=math.inf,
                 bounds=(-math.inf, math.inf), log_images_every=0, compute_diversity=None, group=None, index_offset=0,
                 use_graph=True):
        self.plot_training_dir = plot_training_dir
        self.engine = None
        self.w_inertia = w_inertia
        self.w_cogn = w_cogn
        self.w_soci = w_soci
        self.tolerance = tolerance
        self.n_iterations = n_iterations
        self.dim_space = dim_space
        self.num_particles = num_particles
        self.discovery = discovery
        self.device = torch.device(device)
        self.seed, self.v_max, self.bounds = seed, v_max, bounds
        self.log_images_every = log_images_every
        self.group, self.index_offset, self.use_graph = group, index_offset, use_graph
        hist_bytes = 2 * (n_iterations + 1) * num_particles * dim_space * 4
        self.record_history = (hist_bytes <= (1 << 30)) if record_history is None else record_history
        # Swarm.mse is O(P^2 D): exact up to DIVERSITY_EXACT_MAX particles, above that the same statistic on an evenly strided
        # sample of that many particles (an estimate of the mean pairwise distance; history['mean_mse'] is never left empty)
        self.compute_diversity = True if compute_diversity is None else compute_diversity

        self.g_best_pos = np.empty(self.dim_space)
        self.g_best_val = [np.inf]
        self.g_best_val_dummy = []
        self.swarm = []
        self._hist = None
        self._hist_len = 0
        self._targets = None
        self._inertia_decay = 1.0

    DIVERSITY_EXACT_MAX = 8192       # 0.3 ms per call on a B200 (16 384: 1.2 ms, half of a 65 536-particle iteration)

    # -- coefficients: assignments reach the engine (the move kernel reads w from device state) ----------------
    @property
    def w_inertia(self):
        return self._w_inertia

    @w_inertia.setter
    def w_inertia(self, w):
        self._w_inertia = w
        if self.engine is not None:
            self.engine.w0 = float(w)
            self.engine.set_inertia(float(w))

    @property
    def w_cogn(self):
        return self._w_cogn

    @w_cogn.setter
    def w_cogn(self, c):
        self._w_cogn = c
        if self.engine is not None and self.engine.c_cogn != float(c):
            self.engine.c_cogn = float(c)
            self.engine._graph = None              # launch arguments are baked into a captured iteration

    @property
    def w_soci(self):
        return self._w_soci

    @w_soci.setter
    def w_soci(self, c):
        self._w_soci = c
        if self.engine is not None and self.engine.c_soci != float(c):
            self.engine.c_soci = float(c)
            self.engine._graph = None

    # -- engine construction -------------------------------------------------------------------
    def _build_engine(self, targets=None):
        P, D = self.num_particles, self.dim_space
        with torch.cuda.device(self.device):
            gen = self.discovery.generator_net(P)
            fit = self.discovery.assessor_fitness(P)
            if targets is not None:
                fit = _engine.ReconFitness(targets, w_rec=1.0, eps=0.1, assessor=fit, w_ass=1.0)
            self.engine = _engine.SwarmEngine(P, D, generator=gen, fitness=fit, w_inertia=self.w_inertia, w_cogn=self.w_cogn,
                                              w_soci=self.w_soci, v_max=self.v_max, lo=self.bounds[0], hi=self.bounds[1],
                                              seed=self.seed, index_offset=self.index_offset,
                                              max_iterations=max(self.n_iterations, 1), group=self.group, device=self.device,
                                              need_image_f32=bool(self.log_images_every))
        self.engine.inertia_decay = self._inertia_decay
        self.swarm = _ParticleViews(self, P)
        if self.record_history:
            self._hist = torch.empty(self.n_iterations + 1, 2, P, D, dtype=torch.float32, device=self.device)
            self._hist_len = 0

    def _record(self):
        if self._hist is not None and self._hist_len < self._hist.shape[0]:
            self._hist[self._hist_len, 0].copy_(self.engine.pos[:, :self.dim_space])
            self._hist[self._hist_len, 1].copy_(self.engine.vel[:, :self.dim_space])
            self._hist_len += 1

    def _history_of(self, i, which):
        if self._hist is None:
            return []
        return list(self._hist[:self._hist_len, which, i].double().cpu().numpy())

    # -- reference methods ---------------------------------------------------------------------
    def update_inertia(self):
        self.w_inertia = 0.99 * self.w_inertia          # the setter writes the new weight into the engine's device state
        print(f"Update Inertia: {self.w_inertia}")

    def mse(self):
        """Mean Euclidean distance over all unordered particle pairs (src/pso/util_pso.py:76-86); above DIVERSITY_EXACT_MAX
        particles, over the pairs of an evenly strided sample of that size."""
        pos = self.engine.pos[:, :self.dim_space]
        if pos.shape[0] > self.DIVERSITY_EXACT_MAX:
            idx = torch.linspace(0, pos.shape[0] - 1, self.DIVERSITY_EXACT_MAX, device=pos.device).long()
            pos = pos.index_select(0, idx).contiguous()
        return _engine.mean_pairwise_distance(pos)

    def _reset_best(self):
        """A fresh run on the same object (second optimize()): the host-side mirrors restart with the engine's state."""
        self.g_best_pos = np.empty(self.dim_space)
        self.g_best_val = [np.inf]
        self.g_best_val_dummy = []
        self._hist_len = 0

    def init_particles(self, generator: torch.Generator = None, pos=None, vel=None):
        if self.engine is None:
            self._build_engine()
        self._reset_best()
        self.engine.w0 = float(self._w_inertia)
        self.engine.init_particles(pos=pos, vel=vel, generator=generator)
        self._record()
        return self.engine.pos[:, :self.dim_space].double().cpu().numpy()

    def init_particles_pso_inverter(self, ood_patient, encoder):
        """pos = E(img) per image (src/pso/util_pso.py:93-112), batched; velocity random as in :31."""
        imgs = []
        for img_idx, (img, _) in enumerate(ood_patient):
            if img_idx >= self.num_particles:
                break
            assert img.dtype == torch.float32
            assert torch.max(img) <= 1.0
            assert torch.min(img) >= -1.0
            assert img[0].shape[0] == 1
            imgs.append(img[:1].to(self.device))
        self.num_particles = len(imgs)
        targets = torch.cat(imgs, dim=0)
        with torch.no_grad():
            latent = encoder(targets.float()).reshape(len(imgs), -1)
        self._targets = targets
        self._build_engine(targets=targets.reshape(len(imgs), -1))
        self.swarm._imgs = dict(enumerate(imgs))
        self._reset_best()
        self.engine.init_particles(pos=latent)
        self._record()
        return self.engine.pos[:, :self.dim_space].double().cpu().numpy()

    def update_personal_best(self, iteration, ood_patient=None):
        """Generator + fitness for the whole swarm, then the strict-< personal best (src/pso/util_pso.py:114-124), as the
        engine's fused launches (networks + psob_fitness_tail)."""
        self.engine.evaluate_and_personal_best()
        if self.log_images_every and iteration % self.log_images_every == 0:
            self._dump_images(iteration)
        self._pending_best = True

    def update_global_best(self):
        if not getattr(self, "_pending_best", False):
            raise _ffi.PsobError("update_global_best() follows update_personal_best(): the min-loc key is produced there")
        self.engine.global_best()
        self._pending_best = False
        self._sync_global_best()

    def _sync_global_best(self):
        """Mirror the device-side global best into the reference's host attributes: one packed device -> host copy."""
        e = self.engine
        first = len(self.g_best_val_dummy)
        it, w, vals, replaced, gpos = e.report(first)
        for k in range(len(vals)):
            v = np.asarray([vals[k]], dtype=np.float32)
            if replaced[k]:
                if len(self.g_best_val) == 1 and np.all(np.isinf(self.g_best_val[-1])):
                    self.g_best_val[-1] = v       # first improvement overwrites the inf (src/pso/util_pso.py:146-147)
                else:
                    self.g_best_val.append(v)
            self.g_best_val_dummy.append(v)
        self.g_best_pos = gpos[:self.dim_space].astype(np.float64)
        if self._inertia_decay != 1.0:
            self._w_inertia = w                    # mirror of the device-side schedule (no write-back)

    def move_particles(self):
        self.engine.move()
        self._record()

    def checkpoint(self):
        history_particles, history_particles_vel = {}, {}
        if self._hist is None:
            return history_particles, history_particles_vel
        h = self._hist[:self._hist_len].double().cpu().numpy()
        for p_idx in range(self.num_particles):
            history_particles[f"particle_{p_idx}"] = pd.DataFrame(h[:, 0, p_idx])
            history_particles_vel[f"particle_{p_idx}"] = pd.DataFrame(h[:, 1, p_idx])
        return history_particles, history_particles_vel

    def _dump_images(self, iteration):
        import torchvision
        resc = torch.empty_like(self.engine.gen.image())
        _engine.rescale_minmax(self.engine.gen.image(), resc)
        grid = torchvision.utils.make_grid(resc.view(-1, 1, 28, 28)[:256].cpu())
        os.makedirs(self.plot_training_dir, exist_ok=True)
        torchvision.utils.save_image(grid, os.path.join(self.plot_training_dir, f"pso_images_{iteration}.png"))

    def optimize(self, schedule_inertia=False, early_stopping=True, ood_patient=None, encoder=None):
        self._inertia_decay = 0.99 if schedule_inertia else 1.0
        if ood_patient is not None:
            self.init_particles_pso_inverter(ood_patient=ood_patient, encoder=encoder)
        else:
            self.init_particles()
        self.engine.inertia_decay = self._inertia_decay
        i = 1
        history = {"mean_mse": [], "global_best_val": []}
        # Host work per iteration (diversity print, early stop, history) needs a sync; without it
        # the iterations run back to back as replays of one CUDA graph.
        per_iter_host = self.compute_diversity or early_stopping or self.record_history or self.log_images_every
        if not per_iter_host:
            self.engine.run(self.n_iterations, use_graph=self.use_graph)
            self._sync_global_best()
            i = self.n_iterations + 1
        else:
            while i < self.n_iterations + 1:
                self.update_personal_best(iteration=i, ood_patient=ood_patient)
                self.update_global_best()
                self.move_particles()
                if self.compute_diversity:
                    d_euc = self.mse()
                    history["mean_mse"].append(d_euc)
                    print(f"Iteration: {i}, mean euclidean distance: {d_euc}, global best value: {self.g_best_val[-1]}")
                if i > 2 and len(self.g_best_val) > 2 and early_stopping:
                    if abs(self.g_best_val[-1] - self.g_best_val[-2]) < self.tolerance:
                        break
                i += 1
        history_particles, history_particles_vel = self.checkpoint()
        print(f"The best position is: {self.g_best_pos} with value: {self.g_best_val[-1]}, in iteration number: {i}")
        return history, history_particles, history_particles_vel, i
