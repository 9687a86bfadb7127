"""Generic-objective swarm with the tutorial's interface (src/hands_on/pso.py:24-87) on the
B200 engine: `obj_fun` maps a position tensor [P, D] (cuda) to fitness [P]; a plain python
callable on one position (the tutorial's form) is applied row by row on the host."""
from __future__ import annotations

import numpy as np
import torch

from psob import engine as _engine


class Swarm:
    def __init__(self, obj_fun, num_particles, n_iterations, dim_space, tolerance=0.0001, w_inertia=0.5, w_cogn=0.8,
                 w_soci=0.9, *, batched=False, seed=42, device="cuda:0"):
        self.w_inertia, self.w_cogn, self.w_soci = w_inertia, w_cogn, w_soci
        self.tolerance = tolerance
        self.n_iterations = n_iterations
        self.dim_space = dim_space
        self.g_best_pos = np.empty(self.dim_space)
        self.g_best_val = [np.inf]
        self.num_particles = num_particles
        self.obj_fun = obj_fun
        self.batched = batched
        self.device = torch.device(device)
        self.engine = _engine.SwarmEngine(num_particles, dim_space, objective=self._objective, w_inertia=w_inertia,
                                          w_cogn=w_cogn, w_soci=w_soci, seed=seed, max_iterations=n_iterations,
                                          device=self.device)

    def _objective(self, pos, fit):
        p = pos[:, :self.dim_space]
        if self.batched:
            fit.copy_(self.obj_fun(p).to(torch.float32))
        else:
            host = p.double().cpu().numpy()
            fit.copy_(torch.tensor([float(self.obj_fun(row)) for row in host], dtype=torch.float32))

    def init_particles(self):
        g = torch.Generator(device=self.device).manual_seed(self.engine.seed)
        pos = torch.rand(self.num_particles, self.dim_space, generator=g, device=self.device) * 10 - 5     # U(-5,5), pso.py:6
        vel = (torch.randn(self.num_particles, self.dim_space, generator=g, device=self.device) - 0.5) / 10
        self.engine.init_particles(pos=pos, vel=vel)
        return pos.double().cpu().numpy()

    def optimize(self):
        self.init_particles()
        i = 1
        while i < self.n_iterations + 1:
            self.engine.step()
            vals, replaced = self.engine.history()
            if replaced[-1]:
                if self.g_best_val[-1] == np.inf:
                    self.g_best_val[-1] = float(vals[-1])
                else:
                    self.g_best_val.append(float(vals[-1]))
            if i > 2 and len(self.g_best_val) > 2:
                if abs(self.g_best_val[-1] - self.g_best_val[-2]) < self.tolerance:
                    break
            i += 1
        self.g_best_pos = self.engine.gbest_pos[:self.dim_space].double().cpu().numpy()
        print(f"The best position is: {self.g_best_pos} with value: {self.g_best_val[-1]}, in iteration number: {i}")
