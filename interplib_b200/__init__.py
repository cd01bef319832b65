"""interplib_b200 -- B200-native (sm_100a, FP64) implementation of interplib's element-batched
basis-evaluation / integration hot path behind a flat C ABI (``include/interp_b200.h``).

``interplib_b200.batch``      element-batched NumPy front-end (one call = many elements on the GPU)
``interplib_b200.interp``     drop-in mirror of the reference's per-element API (``interplib._interp`` +
                              ``interplib.integration``): same names, arguments and error behaviour, every
                              numeric call is a batch of one on the GPU.
``interplib_b200.partition``  element partition over the GPUs of one box + the two NCCL collectives (C ABI)
``interplib_b200.synthetic``  seeded inputs of the BASELINE configurations (pure NumPy: importing it does not
                              load ``libinterp_b200.so``)

The shared library is loaded on first use of anything that needs it (PEP 562), so tools that only want the
synthetic inputs -- e.g. the reference arm of ``bench.py`` -- never map it.
"""
import importlib

_LAZY = {
    "batch": (".batch", None),
    "Context": (".batch", "Context"),
    "default_context": (".batch", "default_context"),
    "device_count": (".batch", "device_count"),
    "InterpError": ("._lib", "InterpError"),
    "HOST": ("._lib", "HOST"),
    "DEVICE": ("._lib", "DEVICE"),
    "LIB_PATH": ("._lib", "LIB_PATH"),
}

__all__ = sorted(_LAZY)


def __getattr__(name):
    try:
        module, attr = _LAZY[name]
    except KeyError:
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}") from None
    mod = importlib.import_module(module, __name__)
    value = mod if attr is None else getattr(mod, attr)
    globals()[name] = value
    return value
