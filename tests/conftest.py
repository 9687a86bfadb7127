import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gan-discovery-pso_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    path = os.path.join(ROOT, "tests", "golden", "ref_golden.npz")
    with np.load(path) as z:
        return {k: z[k] for k in z.files}


def load_sd(golden, prefix):
    import torch
    pre = prefix + "/"
    return {k[len(pre):]: torch.from_numpy(v.copy()) for k, v in golden.items()
            if k.startswith(pre) and not k.endswith("/out") and k[len(pre):].split(".")[0] in
            ("gen", "disc", "enc", "decoder", "0", "1")}


@pytest.fixture(scope="session")
def golden_resnet():
    """Outputs of the reference's own `Resnet` (src/pso/util_cnn.py) on seeded parameters (oracle/gen_golden_resnet.py)."""
    path = os.path.join(ROOT, "tests", "golden", "ref_resnet.npz")
    with np.load(path) as z:
        return {k: z[k] for k in z.files}
