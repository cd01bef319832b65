"""Drop-in acceptance, host part: the reference's own test files that only need the table layer (rules, Lagrange /
Bernstein polynomials, FunctionSpace.evaluate, CovectorBasis algebra) run UNCHANGED against this repo's ``interplib``
alias, in both layouts (tests/_dropin.py).  The files whose numeric calls run on the GPU are in test_dropin_gpu.py."""
import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent))
import _dropin  # noqa: E402

pytestmark = pytest.mark.skipif(not _dropin.available(), reason="oracle/_ref/tests missing (run __graft_entry__.build() "
                                                               "where /root/reference is mounted)")


@pytest.mark.parametrize("layout", ["package", "extension"])
def test_reference_host_tests_pass_unchanged(tmp_path, layout):
    res = _dropin.run(layout, _dropin.HOST_FILES, tmp_path)
    assert res["returncode"] == 0, res["stdout"] + res["stderr"]
    assert res["counts"].get("passed", 0) >= 250 and not res["counts"].get("failed"), res["summary"]


def test_alias_exports_every_name_the_reference_package_does():
    """python/interplib/__init__.py:3-50 of the reference: every public name resolves in the alias package."""
    import ast

    ref_init = _dropin.REF_PKG / "__init__.py"
    names = set()
    for node in ast.walk(ast.parse(ref_init.read_text())):
        if isinstance(node, ast.ImportFrom):
            names.update(a.asname or a.name for a in node.names)
    mine = set()
    for node in ast.walk(ast.parse((_dropin.ROOT / "dropin" / "interplib" / "__init__.py").read_text())):
        if isinstance(node, ast.ImportFrom):
            mine.update(a.asname or a.name for a in node.names)
    assert names <= mine, sorted(names - mine)
