"""A/B tuning knobs.  PSOB_* environment variables are honoured only when PSOB_TUNING=1, so that a stray variable in a
user's shell never changes how the product path executes (csrc/pixel_gemm.cu applies the same gate to the native knobs).
tools/ and the mode-equivalence tests set PSOB_TUNING=1 explicitly."""
import os


def tuning_enabled() -> bool:
    return os.environ.get("PSOB_TUNING") == "1"


def knob(name: str, default: str) -> str:
    return os.environ.get(name, default) if tuning_enabled() else default
