"""xpano_b200: B200-native full-resolution export hot path of krupkat/xpano
(warp -> seams/masks -> multiband blend) behind the reference's operator
surface. The compute lives in libxpano_b200.so (hand-written sm_100a CUDA
behind a C ABI, include/xpano_b200.h); these modules are thin ctypes mirrors
of the reference interfaces. No CPU fallback."""
from . import _lib  # noqa: F401
from . import sharding  # noqa: F401
from .auto_crop import FindLargestCrop  # noqa: F401
from .blenders import Multiblend, MultiBandOpenCV  # noqa: F401
from .codec import JpegCodec  # noqa: F401
from .exposure import (BlocksGainCompensator, ComputeSeamScale,  # noqa
                       EstimateSeams)
from .stitcher import Composer, compose_panorama  # noqa: F401
from .warpers import (BORDER_CONSTANT, BORDER_REFLECT, INTER_LINEAR,  # noqa
                      INTER_NEAREST, PickWarper, PickWarperPortrait,
                      ProjectionType, Rotate, RotationWarper, WarperCreator)
