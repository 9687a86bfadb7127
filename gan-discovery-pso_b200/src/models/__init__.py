"""`src.models` is empty in the reference (placeholder for a missing StyleGAN submodule);
north_star names it, so the drop-in model classes are re-exported here."""
from src.utils.util_dcgan import Discriminator, Encoder, Generator, initialize_weights  # noqa: F401
