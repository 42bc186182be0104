"""B200-native (sm_100a) sparse-convolution-noise GPIS medium query — drop-in for that one path of
PeterZs/sparse-conv-gpis-tungsten.  See DESIGN.md; the C ABI is include/scn.h."""
from . import abi  # noqa: F401
from .medium import MediumFactory, SparseConvolutionNoiseMedium, new_state  # noqa: F401

__all__ = ["abi", "MediumFactory", "SparseConvolutionNoiseMedium", "new_state"]
