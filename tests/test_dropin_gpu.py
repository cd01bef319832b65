"""Drop-in acceptance on the B200: the reference's own test files for the
hot path -- mass matrices, L2 projection, DoF reconstruction, coordinate / space maps, gradient mass matrices,
incidence -- run UNCHANGED with ``import interplib`` resolved to this repository, in both layouts of tests/_dropin.py.
Every numeric call in them goes through libinterp_b200.so on the GPU (batches of one element)."""
import json
import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent))
import _dropin  # noqa: E402

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not _dropin.available(), reason="oracle/_ref/tests missing (built by __graft_entry__.build() "
                                                                 "where /root/reference is mounted; travels with gpurun)")]


@pytest.mark.parametrize("layout", ["package", "extension"])
def test_reference_hot_path_tests_pass_unchanged(tmp_path, layout):
    res = _dropin.run(layout, _dropin.GPU_FILES + _dropin.HOST_FILES, tmp_path)
    out = _dropin.ROOT / "gpurun_out"
    if out.is_dir():
        (out / f"dropin_{layout}.json").write_text(json.dumps({k: res[k] for k in ("returncode", "summary", "counts")}))
        (out / f"dropin_{layout}.log").write_text(res["stdout"] + "\n---- stderr\n" + res["stderr"])
    assert res["returncode"] == 0, res["stdout"] + res["stderr"]
    assert res["counts"].get("passed", 0) >= 1200 and not res["counts"].get("failed"), res["summary"]
