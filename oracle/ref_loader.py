"""TEST INFRASTRUCTURE — loader for the *real* reference classes (container only).

The reference (`/root/reference`, read-only, absent on the GPU box) is a pure
Python/PyTorch repo.  Its PSO path imports plotting/reporting modules that are
not installed here (matplotlib, natsort, openpyxl, torchsummary, plotly,
seaborn, umap, pydicom); none of them carries arithmetic, so
they are replaced by MagicMock before import.  Nothing under `oracle/` is ever
imported by the product (`gan-discovery-pso_b200/`); only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg use it, and this
file in particular is used only by `oracle/gen_golden.py` and by the
container-only tests that pin `oracle/port.py` against the live reference.
"""
import os
import sys
import types
from unittest import mock

REFERENCE_ROOT = os.environ.get("PSOB_REFERENCE_ROOT", "/root/reference")

_STUBS = [
    "matplotlib", "matplotlib.pyplot", "matplotlib.cm", "matplotlib.ticker",
    "matplotlib.colors", "matplotlib.patches", "matplotlib.animation",
    "natsort", "openpyxl", "torchsummary", "plotly", "plotly.express",
    "plotly.graph_objects", "seaborn", "umap", "pydicom", "imageio",
]


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "pso"))


def load():
    """Import the reference modules on the PSO path; returns a namespace."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    for name in _STUBS:
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = mock.MagicMock(name=name)
    # The product mirrors the reference's `src.*` import paths; make sure the
    # reference's own tree wins inside this loader by importing under a
    # private alias namespace.
    saved = {k: v for k, v in sys.modules.items() if k == "src" or k.startswith("src.")}
    for k in saved:
        del sys.modules[k]
    # the reference's `src` has no __init__.py (a namespace package): any REGULAR package called `src` further down sys.path --
    # the product's drop-in mirror -- would win the import, so such entries are hidden while the reference loads
    old_path = list(sys.path)
    sys.path[:] = [q for q in old_path if not os.path.isfile(os.path.join(q or ".", "src", "__init__.py"))]
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        import src.pso.util_pso as util_pso
        import src.pso.util_discovery as util_discovery
        import src.utils.util_dcgan as util_dcgan
        import src.utils.util_data as util_data
        import src.inverter.utils_ae.util_inverter as util_inverter
        import src.inverter.utils_vq_vae.util_model as vq_model
        import src.inverter.utils_vq_vae.util_function as vq_function
        import src.pso.util_cnn as util_cnn
        sys.path.insert(0, os.path.join(REFERENCE_ROOT, "src"))
        try:
            import hands_on.pso as hands_on_pso
        finally:
            sys.path.remove(os.path.join(REFERENCE_ROOT, "src"))
    finally:
        sys.path[:] = old_path
        ref_mods = {k: v for k, v in sys.modules.items() if k == "src" or k.startswith("src.")}
        for k in ref_mods:
            del sys.modules[k]
        sys.modules.pop("hands_on", None)
        sys.modules.pop("hands_on.pso", None)
        sys.modules.update(saved)
    # figure rendering in Swarm.update_personal_best is I/O only: neutralise.
    util_pso.show = lambda imgs: None
    util_pso.plt = mock.MagicMock(name="plt")
    ns = types.SimpleNamespace(
        util_pso=util_pso, util_discovery=util_discovery, util_dcgan=util_dcgan,
        util_data=util_data, util_inverter=util_inverter, vq_model=vq_model,
        vq_function=vq_function, hands_on_pso=hands_on_pso, util_cnn=util_cnn)
    return ns
