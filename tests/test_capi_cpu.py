"""CPU: the C-ABI library loads, exports every symbol include/interp_b200.h declares, and refuses -- loudly --
to run the batched hot path without a GPU (no CPU fallback)."""
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared_symbols():
    text = (ROOT / "include" / "interp_b200.h").read_text()
    return sorted(set(re.findall(r"IB200_API\s+[\w\s\*]+?\b(ib200_\w+)\s*\(", text)))


def test_header_symbols_exported():
    from interplib_b200 import _lib

    names = declared_symbols()
    assert len(names) >= 35
    out = subprocess.run(["nm", "-D", "--defined-only", str(_lib.LIB_PATH)], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\sT\s+(ib200_\w+)", out))
    missing = [n for n in names if n not in exported]
    assert not missing, f"declared in the header but not exported: {missing}"
    # the ctypes prototypes cover the same set
    assert sorted(_lib.EXPORTED_SYMBOLS) == names
    # nothing but the C ABI leaks out of the library
    leaked = [s for s in re.findall(r"\s[TDB]\s+(\w+)", out) if not s.startswith("ib200_")]
    assert not leaked, leaked


def test_library_is_sm100a_cuda():
    from interplib_b200 import _lib

    out = subprocess.run(["cuobjdump", "-lelf", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_error_strings():
    from interplib_b200._lib import lib

    assert lib.ib200_error_str(0) == b"INTERP_SUCCESS"
    assert lib.ib200_error_str(5) == b"INTERP_ERROR_INVALID_ENUM"  # same numbering as src/common/error.h:9-25
    assert lib.ib200_error_str(11) == b"INTERP_ERROR_MATRIX_DIMS_MISMATCH"
    assert lib.ib200_error_str(101) == b"IB200_ERROR_NO_DEVICE"


def test_invalid_enum_is_value_error():
    from interplib_b200 import batch

    with pytest.raises(ValueError):
        batch.rule_1d(7, 3)
    with pytest.raises(ValueError):
        batch.basis_table_1d((0, 3), ("gauss", 3))


def test_no_cpu_fallback():
    import interplib_b200 as ib

    if ib.device_count() > 0:
        pytest.skip("a GPU is visible here")
    with pytest.raises(ib.InterpError) as e:
        ib.Context(0)
    assert e.value.name == "IB200_ERROR_NO_DEVICE"


def test_product_does_not_touch_oracle():
    """The product path must never import, link or load anything under oracle/."""
    for p in (ROOT / "interplib_b200").rglob("*"):
        if p.suffix in {".py", ".c", ".cu", ".cuh", ".h"} or p.name == "Makefile":
            text = p.read_text()
            assert "oracle" not in text.lower().replace("oracle/_ref", "").replace("the oracle", "") or p.name == "synthetic.py", p
    from interplib_b200 import _lib

    out = subprocess.run(["ldd", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    assert "oracle" not in out
