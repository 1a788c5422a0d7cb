"""Timing of the alignment consumer on resident scores: kernel (CUDA events inside the library) and whole call.
usage: python profiles/time_align.py [segments] [frames_per_segment] [emissions_per_automaton] [columns of the score block]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rasr_b200 import capi  # noqa: E402
from rasr_b200.align import AlignmentAutomata, linear_hmm  # noqa: E402

segs = int(sys.argv[1]) if len(sys.argv) > 1 else 512
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
n = int(sys.argv[3]) if len(sys.argv) > 3 else 300
M = int(sys.argv[4]) if len(sys.argv) > 4 else 4501
eng = capi.Engine(0)
rng = np.random.default_rng(0)
au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, M, n), 3.0, 0.0, 30.0) for _ in range(segs)])
fo = np.arange(segs + 1, dtype=np.uint32) * T
scores = torch.rand((segs * T, M), device="cuda") * 80
kernels = ("ranked",) if os.environ.get("ALIGN_RANKED_ONLY") else ("ranked", "generic")
for name in kernels:
    if name == "generic":
        os.environ["GMM_ALIGN_KERNEL"] = "generic"
    eng.align_device(fo, scores, au, ld=M, want_states=False)
    eng.kernel_timing(True)
    eng.kernel_time()
    t0 = time.perf_counter()
    reps = 5 if name == "ranked" else 1
    for _ in range(reps):
        eng.align_device(fo, scores, au, ld=M, want_states=False)
    wall = (time.perf_counter() - t0) / reps
    ms, k = eng.kernel_time()
    eng.kernel_timing(False)
    print(f"{name}: {segs} segments x {T} frames, {n + 1} states: kernel {ms / k:.3f} ms, call {wall * 1e3:.3f} ms, "
          f"{segs * T / (ms / k) / 1e3:.1f} M frames/s kernel, ordered segments {eng.align_ordered_count}")
