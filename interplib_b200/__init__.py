"""interplib_b200 -- B200-native (sm_100a, FP64) implementation of interplib's element-batched
basis-evaluation / integration hot path behind a flat C ABI (``include/interp_b200.h``).

``interplib_b200.batch``   element-batched NumPy front-end (one call = many elements on the GPU)
``interplib_b200.interp``  drop-in mirror of the reference's per-element API (``interplib._interp`` +
                           ``interplib.integration``): same names, arguments and error behaviour, every
                           numeric call is a batch of one on the GPU.
"""
from . import batch  # noqa: F401
from ._lib import DEVICE, HOST, InterpError, LIB_PATH  # noqa: F401
from .batch import Context, default_context, device_count  # noqa: F401

__all__ = ["batch", "Context", "default_context", "device_count", "InterpError", "HOST", "DEVICE", "LIB_PATH"]
