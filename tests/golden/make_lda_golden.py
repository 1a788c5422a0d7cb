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

# lda_window.npz: the whole front end as share/lda.flow wires it -- the reference's own sliding-window node
# (oracle/_ref/ref_window: Signal::DelayNode through signal-vector-f32-sequence-concatenation in a Flow::Network,
# margin-condition present-not-empty, margin-policy copy) followed by the reference's own matrix product
# (oracle/_ref/ref_matmul), per segment.  <case>_base [frames x baseDim], _frame_offsets, _length, _right, _matrix,
# _window [frames x baseDim*length] and _out [frames x D] as the reference produced them.
assert O.ref_window_available()
rng = np.random.default_rng(2025)
cases = {}
for name, (base_dim, length, right, D, lens) in {
        "lda16x9": (16, 9, 4, 45, [40, 1, 2, 9, 3, 120]),          # 9-frame window centred (the usual 16x9 -> 45)
        "causal7x3": (7, 3, 0, 5, [1, 2, 3, 30]),                  # right = 0: past frames only
        "lookahead5x6": (5, 6, 5, 8, [4, 6, 7, 25]),               # right = length-1: future frames only
        "wide3x17": (3, 17, 8, 11, [5, 16, 17, 18, 60])}.items():  # window longer than some segments
    fo = np.cumsum([0] + lens).astype(np.uint32)
    base = (rng.standard_normal((int(fo[-1]), base_dim)) * 10).astype(np.float32)
    m = rng.standard_normal((D, base_dim * length)).astype(np.float32)
    win = np.concatenate([O.run_ref_window(base[fo[i]:fo[i + 1]], length, right) for i in range(len(lens))])
    assert win.shape == (int(fo[-1]), base_dim * length)            # present-not-empty: one output per input frame
    cases.update({name + "_base": base, name + "_frame_offsets": fo, name + "_length": np.uint32(length),
                  name + "_right": np.uint32(right), name + "_matrix": m, name + "_window": win,
                  name + "_out": O.run_ref_matmul(m, win)})
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "lda_window.npz"), **cases)
print({k: v.shape for k, v in cases.items() if k.endswith(("_window", "_out"))})
