"""TEST / BASELINE INFRASTRUCTURE -- never imported by the product.

Times the UNMODIFIED reference on the host cores: `Swarm.optimize` (src/pso/util_pso.py:167-193) driven exactly as
`src/training/pso_discovery.py:200-219` drives it, BASELINE config 1:

    Generator(100, 1, 64) + initialize_weights                       (src/utils/util_dcgan.py:45-48,128-149)
    get_cnn('ResNet50', 1, iid_classes=[0,2,3,4,6,7,8,9], n_class=8) + initialize_weights(.., 'glorot_normal')
    Discovery(iid_class=0, ...), Swarm(P=64, n_iterations, dim_space=100, tolerance=1e-5, w=0.73, c=1.496/1.496)
    optimize(early_stopping=False), numpy / torch seeds 42

The reference classes come from oracle/_ref (a verbatim copy made by oracle/make_ref.sh; what the GPU box has) or, in the
build container, from /root/reference.  Only I/O is neutralised (oracle/ref_loader.py: plotting modules stubbed, `show`,
`plt`; here also `make_grid`, whose result only feeds a PNG and TensorBoard); every arithmetic line that runs is the
reference's own.  The five methods BASELINE.md names are wrapped with wall-clock timers for the breakdown."""
import contextlib
import io
import os
import tempfile
import time
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
IID_CLASSES = [0, 2, 3, 4, 6, 7, 8, 9]


def reference_root():
    local = os.path.join(HERE, "_ref")
    if os.path.isdir(os.path.join(local, "src", "pso")):
        return local
    if os.path.isdir("/root/reference/src/pso"):
        return "/root/reference"
    return None


def load_reference():
    from oracle import ref_loader
    root = reference_root()
    if root is None:
        raise RuntimeError("no reference tree: run oracle/make_ref.sh in the build container (oracle/_ref travels with gpurun)")
    ref_loader.REFERENCE_ROOT = root
    return ref_loader.load(), root


def time_config1(num_particles=64, n_iterations=50, seed=42, quiet=True):
    """-> dict(evals_per_s, seconds, particles, iterations, breakdown_s{method: seconds}, threads, cpu_count, root, ...)."""
    ref, root = load_reference()
    U, Dm, G, CNN = ref.util_pso, ref.util_discovery, ref.util_dcgan, ref.util_cnn
    torch.manual_seed(seed)
    np.random.seed(seed)
    gen = G.Generator(z_dim=100, channels_img=1, features_g=64)
    G.initialize_weights(gen)
    cnn = CNN.get_cnn("ResNet50", 1, iid_classes=IID_CLASSES, n_class=8)
    CNN.initialize_weights(cnn, "glorot_normal")
    gen.eval()
    cnn.eval()
    # I/O only: the image grid feeds a PNG (plt is stubbed) and TensorBoard
    tv = types.SimpleNamespace(utils=types.SimpleNamespace(make_grid=lambda imgs, *a, **k: torch.zeros(3, 8, 8)))
    U.torchvision = tv
    breakdown = {}

    def wrap(obj, name, key=None):
        fn = getattr(obj, name)

        def timed(*a, **k):
            t0 = time.perf_counter()
            try:
                return fn(*a, **k)
            finally:
                breakdown[key or name] = breakdown.get(key or name, 0.0) + time.perf_counter() - t0
        setattr(obj, name, timed)

    with tempfile.TemporaryDirectory() as td:
        disc = Dm.Discovery(iid_class=0, model_gan=gen, model_cnn=cnn, device="cpu", iid_classes=IID_CLASSES)
        sw = U.Swarm(td, disc, num_particles=num_particles, n_iterations=n_iterations, dim_space=100, device="cpu",
                     tolerance=1e-5, w_inertia=0.73, w_cogn=1.496, w_soci=1.496)
        for name in ("update_personal_best", "move_particles", "mse", "update_global_best"):
            wrap(sw, name)
        wrap(disc, "particles_to_img")
        out = io.StringIO()
        t0 = time.perf_counter()
        with (contextlib.redirect_stdout(out) if quiet else contextlib.nullcontext()):
            _, _, _, it = sw.optimize(schedule_inertia=False, early_stopping=False)
        dt = time.perf_counter() - t0
        gbest = float(np.asarray(sw.g_best_val[-1]).reshape(-1)[0])
    assert it == n_iterations + 1
    evals = num_particles * n_iterations
    return dict(evals_per_s=evals / dt, seconds=dt, particles=num_particles, iterations=n_iterations,
                breakdown_s={k: round(v, 4) for k, v in breakdown.items()}, threads=torch.get_num_threads(),
                cpu_count=os.cpu_count(), root=os.path.relpath(root, os.path.dirname(HERE)) if root.startswith(os.path.dirname(HERE)) else root,
                gbest=gbest, torch=torch.__version__, numpy=np.__version__)


if __name__ == "__main__":
    import json
    import sys
    sys.path.insert(0, os.path.dirname(HERE))
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    print(json.dumps(time_config1(n_iterations=n)))
