"""Regenerates tests/golden/oracle_golden.npz from the oracle (run from the repo root:
`python tests/golden/make_golden.py`).  The reference ships no fixtures
(SURVEY.md section 0); these pin the oracle against later edits of itself."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402
import make_golden_cases as G  # noqa: E402
from oracle import oracle as O  # noqa: E402

out = {}
for name, pb in G.problems().items():
    q, k, v, dout = cases.make_inputs(pb, seed=G.INPUT_SEED)
    f = O.forward(pb, q, k, v)
    b = O.backward(pb, dout, q, k, v, f["out"], f["lse"])
    out[f"{name}.out"] = f["out"]
    out[f"{name}.lse"] = f["lse"]
    for t in ("dq", "dk", "dv"):
        out[f"{name}.{t}"] = b[t]
np.savez_compressed(os.path.join(HERE, "oracle_golden.npz"), **out)
print("wrote", len(out), "arrays")
