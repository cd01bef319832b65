#!/usr/bin/env python3
"""Build the UNMODIFIED reference (j4nr0th/interplib) into ``oracle/_ref``.

TEST INFRASTRUCTURE ONLY -- nothing under ``oracle/`` is on the product path.

The reference's own build (CMake + scikit-build, ``/root/reference/CMakeLists.txt``) cannot
run because its ``cutl`` submodule is absent (``/root/reference/.SUBMODULES.json``); the only
things it needs from it are an allocator vtable and a multi-dimensional index iterator, which
``oracle/cutl_shim`` provides (no arithmetic).  This script compiles the reference's C sources
*where they lie* under ``/root/reference`` with plain ``gcc`` and puts the outputs only into
``oracle/_ref`` (git-ignored, but shipped to the GPU box):

  oracle/_ref/interplib/_interp.<EXT_SUFFIX>   the reference CPython extension
  oracle/_ref/interplib/*.py                    install-time copy of python/interplib
  oracle/_ref/tests/*.py                        install-time copy of the reference's tests (run unchanged
                                                against this repo's `interplib` alias, tests/test_dropin_gpu.py)
  oracle/_ref/libinterp_ref.so                  the reference C core (hidden symbols made
                                                visible with -DINTERPLIB_INTERNAL=) for ctypes

Flags follow the reference (``CMakeLists.txt:96-102,130-136``): C17 (gnu17 because ``M_PI`` is
used), ``-O2``, ``INTERPLIB_ASSERTS`` / ``CPYUTL_ENABLE_ASSERTS``.  No ``-march`` flag is given so
that no FMA contraction happens: results are the plain IEEE double results the upstream wheel
produces on x86-64.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path(os.environ.get("INTERPLIB_REFERENCE", "/root/reference"))
OUT = HERE / "_ref"
OBJ = OUT / "obj"
SHIM = HERE / "cutl_shim"

CORE_DIRS = ["common", "polynomials", "integration", "basis", "operations", "topology"]


def _run(cmd: list[str]) -> None:
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
        raise RuntimeError(f"command failed: {cmd[0]} ... {cmd[-1]}")


def _compile(src: Path, obj: Path, flags: list[str]) -> Path:
    obj.parent.mkdir(parents=True, exist_ok=True)
    if obj.exists() and obj.stat().st_mtime > src.stat().st_mtime:
        return obj
    _run(["gcc", "-std=gnu17", "-O2", "-fPIC", "-fopenmp", "-w", *flags, "-c", str(src), "-o", str(obj)])
    return obj


def available() -> bool:
    """True when the reference sources are present (build container only)."""
    return (REF / "src" / "python" / "module.c").exists()


def built() -> bool:
    ext = OUT / "interplib" / ("_interp" + sysconfig.get_config_var("EXT_SUFFIX"))
    return ext.exists() and (OUT / "libinterp_ref.so").exists()


def build(force: bool = False) -> Path:
    if not available():
        if built():
            return OUT
        raise FileNotFoundError(f"reference sources not found under {REF}")
    if force and OUT.exists():
        shutil.rmtree(OUT)
    import numpy as np

    core_src = [p for d in CORE_DIRS for p in sorted((REF / "src" / d).glob("*.c"))]
    py_src = sorted((REF / "src" / "python").glob("*.c")) + sorted((REF / "src" / "python" / "topology").glob("*.c"))
    cpyutl_src = [REF / "cpyutl" / "src" / "cpyutl.c"]

    core_flags = ["-DINTERPLIB_ASSERTS", f"-I{SHIM}"]
    py_flags = core_flags + [
        "-DCPYUTL_ENABLE_ASSERTS",
        f"-I{sysconfig.get_paths()['include']}",
        f"-I{np.get_include()}",
        f"-I{REF / 'cpyutl' / 'src'}",
    ]
    # same core sources, symbols left visible: lets ctypes call the reference's own table code
    vis_flags = core_flags + ["-DINTERPLIB_INTERNAL="]

    jobs = []
    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as pool:
        for s in core_src:
            rel = s.relative_to(REF).with_suffix(".o")
            jobs.append(("core", pool.submit(_compile, s, OBJ / "core" / rel, core_flags)))
            jobs.append(("vis", pool.submit(_compile, s, OBJ / "vis" / rel, vis_flags)))
        for s in py_src + cpyutl_src:
            rel = s.relative_to(REF).with_suffix(".o")
            jobs.append(("py", pool.submit(_compile, s, OBJ / "py" / rel, py_flags)))
        objs: dict[str, list[str]] = {"core": [], "vis": [], "py": []}
        for kind, fut in jobs:
            objs[kind].append(str(fut.result()))

    pkg = OUT / "interplib"
    pkg.mkdir(parents=True, exist_ok=True)
    ext = pkg / ("_interp" + sysconfig.get_config_var("EXT_SUFFIX"))
    _run(["gcc", "-shared", "-fopenmp", *objs["py"], *objs["core"], "-o", str(ext), "-lm"])
    _run(["gcc", "-shared", "-fopenmp", *objs["vis"], "-o", str(OUT / "libinterp_ref.so"), "-lm"])
    for f in (REF / "python" / "interplib").glob("*.py*"):
        shutil.copy2(f, pkg / f.name)
    # install-time copy of the reference's own test-suite: the drop-in acceptance test (tests/test_dropin_gpu.py) runs
    # these files UNCHANGED against the `interplib` alias of this repo on the GPU box, where /root/reference is absent
    tests = OUT / "tests"
    tests.mkdir(parents=True, exist_ok=True)
    for f in (REF / "tests").glob("*.py"):
        shutil.copy2(f, tests / f.name)
    return OUT


def import_reference():
    """Import the compiled reference package (``interplib``) from oracle/_ref."""
    if not built():
        build()
    p = str(OUT)
    if p not in sys.path:
        sys.path.insert(0, p)
    import interplib  # noqa: PLC0415

    return interplib


if __name__ == "__main__":
    out = build(force="--force" in sys.argv)
    print(f"reference built into {out}")
    if "--selftest" in sys.argv:
        env = dict(os.environ, PYTHONPATH=str(out))
        sys.exit(
            subprocess.call(
                [sys.executable, "-m", "pytest", str(REF / "tests"), "-q", "-o", "addopts=", "-p", "no:cacheprovider"],
                env=env,
            )
        )
