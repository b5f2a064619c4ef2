"""B200-native drop-in for the reference package `VisionTransformer` (same sub-modules: vit, transformer, config).

    from VisionTransformer import vit, config          # what the reference's src/train.py:21 does

Every forward/backward runs hand-written sm_100a kernels from libvitb200.so through a C ABI; CPU tensors are refused.
"""
from . import config, transformer, vit  # noqa: F401
