"""Generates tests/golden/lda_matmul.npz from the reference's own Math::Matrix<f32> * Math::Vector<f32>
(oracle/_ref/ref_matmul, built from /root/reference by oracle/ref_build/Makefile).  Run in the build container:
    python tests/golden/make_lda_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

O.build_ref()
rng = np.random.default_rng(2024)
cases = {}
for name, (rows, cols, n) in {"lda45x144": (45, 144, 64), "odd7x5": (7, 5, 9), "wide3x1000": (3, 1000, 4)}.items():
    m = (rng.standard_normal((rows, cols)) * rng.uniform(0.01, 3, (rows, 1))).astype(np.float32)
    v = (rng.standard_normal((n, cols)) * 20).astype(np.float32)
    cases[name + "_matrix"], cases[name + "_in"], cases[name + "_out"] = m, v, O.run_ref_matmul(m, v)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "lda_matmul.npz"), **cases)
print({k: v.shape for k, v in cases.items()})
