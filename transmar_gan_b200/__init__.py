"""transmar_gan_b200 — B200-native (sm_100a) implementation of TransMAR-GAN's NGswin hot path.

Public surface:
    NGswin            drop-in for models.generator.ngswin.NGswin
    ops               torch.ops.b200mar.* custom ops over the C ABI (include/b200mar.h)
    build_library()   compile libb200mar.so in-tree with nvcc for sm_100a
"""
from ._cabi import build as build_library, LIB_PATH  # noqa: F401
from . import ops  # noqa: F401
from .ngswin import NGswin  # noqa: F401

__all__ = ["NGswin", "ops", "build_library", "LIB_PATH"]
