"""The inverter's import path (src/inverter/utils_ae/util_inverter.py:119-184): `Encoder`, `Generator` and
`Discriminator` with the reference's constructors and state_dict keys, running on the B200 engine.  The reference keeps
verbatim copies of the DCGAN modules here; this mirror re-exports the engine-backed ones.  Auto-encoder training
(`train_autoencoder` etc.) is outside the swarm-evaluation scope."""
from src.utils.util_dcgan import Discriminator, Encoder, Generator, initialize_weights  # noqa: F401

__all__ = ["Encoder", "Generator", "Discriminator", "initialize_weights"]
