"""Drop-in import path of the reference (`models/generator/ngswin.py`): every reference script
does `from models.generator.ngswin import NGswin`; here that name resolves to the B200-native
implementation."""
from transmar_gan_b200.ngswin import NGswin  # noqa: F401

__all__ = ["NGswin"]
