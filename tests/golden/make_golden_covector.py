#!/usr/bin/env python3
"""Generate tests/golden/covector.npz -- exhaustive integer tables of the compiled reference's CovectorBasis type
(wedge, Hodge, ordering) for 1..5 dimensions.  Build container only."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import refdrive as rd  # noqa: E402

data = {}
for n in range(1, 6):
    t = rd.covector_table(n)
    for k in ("wedge", "hodge", "order"):
        data[f"n{n}_{k}"] = t[k]
np.savez(Path(__file__).resolve().parent / "covector.npz", **data)
print({k: v.shape for k, v in data.items()})
