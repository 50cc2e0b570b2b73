"""enarf-gan_b200: B200-native (sm_100a) articulated tri-plane radiance-field renderer hot path.

Sub-modules are imported lazily by name (``enarf_gan_b200.rendering`` ...); importing the package
itself never needs the CUDA library, so the oracle/CPU tests can use ``synthetic`` on any box.
"""
__version__ = "0.1.0"
