"""Shared driver of the drop-in acceptance tests: runs the REFERENCE's own test files, unchanged (the install-time copy
oracle/build_ref.py places under oracle/_ref/tests), in a fresh interpreter whose ``import interplib`` resolves to this
repository's implementation.  Two layouts:

``package``    <repo>/dropin on the path: ``interplib`` and all its sub-modules are this repo's (what a user gets by
               putting that directory ahead of an installed reference);
``extension``  only ``interplib._interp`` -- the CPython extension of the reference -- is replaced by this repo's
               stand-in; ``__init__.py``, ``integration.py``, ``domains.py``, ``degrees_of_freedom.py``, ``lagrange.py``,
               ``enum_type.py`` are the reference's own files (install-time copy oracle/_ref/interplib/*.py), running
               unchanged on top of it.  This is the boundary BASELINE.json's north star names.
"""
import os
import re
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
REF_TESTS = ROOT / "oracle" / "_ref" / "tests"
REF_PKG = ROOT / "oracle" / "_ref" / "interplib"

# files that touch only the host table layer (no CUDA device needed) / files whose numeric calls run on the GPU;
# test_base_geo / test_manifold2d / test_mesh1d exercise the topology types, which SURVEY section 2 marks out of scope
HOST_FILES = ["test_lagrange.py", "test_integration_rule.py", "test_integration_space.py", "test_covector_basis.py",
              "test_function_space.py"]
GPU_FILES = ["test_mass_matrix.py", "test_l2_proj.py", "test_dofs.py", "test_mapping.py", "test_gradient.py",
             "test_incidence.py"]


def available() -> bool:
    return REF_TESTS.is_dir() and any(REF_TESTS.glob("test_*.py"))


def make_layout(kind: str, tmp: Path) -> Path:
    """Directory to put first on PYTHONPATH."""
    if kind == "package":
        return ROOT / "dropin"
    if kind == "reference":  # the compiled reference itself (oracle/_ref): the oracle's self-test, nothing of this repo
        return ROOT / "oracle" / "_ref"
    assert kind == "extension"
    pkg = tmp / "ext_layout" / "interplib"
    pkg.mkdir(parents=True, exist_ok=True)
    for f in REF_PKG.glob("*.py"):
        shutil.copy2(f, pkg / f.name)  # the reference's Python layer, byte for byte
    shutil.copy2(ROOT / "dropin" / "interplib" / "_interp.py", pkg / "_interp.py")
    (pkg / "_interp_repo_root.txt").write_text(str(ROOT))
    return pkg.parent


def run(kind: str, files, tmp: Path, timeout: int = 1500) -> dict:
    path = make_layout(kind, tmp)
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([str(path), str(ROOT)])
    env["INTERP_B200_REPO"] = str(ROOT)
    cmd = [sys.executable, "-m", "pytest", "-q", "-o", "addopts=", "-p", "no:cacheprovider", "--rootdir", str(tmp),
           "-W", "ignore::DeprecationWarning", *[str(REF_TESTS / f) for f in files]]
    proc = subprocess.run(cmd, cwd=tmp, env=env, capture_output=True, text=True, timeout=timeout)
    tail = proc.stdout.strip().splitlines()[-1] if proc.stdout.strip() else ""
    counts = {k: int(v) for v, k in re.findall(r"(\d+) (passed|failed|error|errors|skipped|xfailed)", tail)}
    return {"returncode": proc.returncode, "summary": tail, "counts": counts, "stdout": proc.stdout[-6000:],
            "stderr": proc.stderr[-2000:]}
