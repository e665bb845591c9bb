"""B200-native (sm_100a) caption-decoder hot path of the Transform-and-Tell derivative
harsh157/news_image_caption: hand-written CUDA kernels behind a C ABI (include/tnt_b200.h) and a
host-side mirror of the reference's decoder / loss module interface.  No CPU fallback."""
from . import _lib  # noqa: F401
from ._lib import TntError, lib, require_device  # noqa: F401
from .decoder import (AdaptiveLoss, DynamicConvDecoder, DynamicConvDecoderEntityPointer,  # noqa: F401
                      DynamicConvDecoderPointer)
from . import modules, transformer_v2  # noqa: F401  (module-level mirrors; nothing runs at import)

__all__ = ['DynamicConvDecoder', 'DynamicConvDecoderPointer', 'DynamicConvDecoderEntityPointer', 'AdaptiveLoss', 'lib', 'require_device', 'TntError']
