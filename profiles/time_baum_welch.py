"""Timing of the Baum-Welch consumer: kernel on resident (dense) scores, and the features-in training pass.
usage: python profiles/time_baum_welch.py [segments] [frames_per_segment] [emissions_per_automaton]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rasr_b200 import capi, synth, tables  # noqa: E402
from rasr_b200.align import AlignmentAutomata, linear_hmm  # noqa: E402

segs = int(sys.argv[1]) if len(sys.argv) > 1 else 296
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
n = int(sys.argv[3]) if len(sys.argv) > 3 else 300
ms = synth.make_model(M=4501, K=16, D=45, seed=1001)
eng = capi.Engine(0)
t = tables.float_tables(ms)
eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX,
                 kernel=capi.KERNEL_AUTO)
rng = np.random.default_rng(0)
au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, 320, n), 3.0, 0.0, 30.0) for _ in range(segs)])
fo = np.arange(segs + 1, dtype=np.uint32) * T
scores = torch.rand((segs * T, 320), device="cuda") * 80
w = torch.empty((segs * T, 320), device="cuda")
eng.baum_welch_device(fo, scores, au, w)
eng.kernel_timing(True)
eng.kernel_time()
reps = 5
for _ in range(reps):
    eng.baum_welch_device(fo, scores, au, w)
kms, k = eng.kernel_time()
eng.kernel_timing(False)
print(f"bw_kernel: {segs} segments x {T} frames, {n + 1} states, dense 320-column block: {kms / k:.3f} ms per launch = "
      f"{segs * T / (kms / k) / 1e3:.1f} M frames/s")
# features in, estimators updated: compact scoring -> forward-backward -> weighted E-step, two lanes
au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, ms.M, n), 3.0, 0.0, 30.0) for _ in range(segs * 4)])
fo = np.arange(segs * 4 + 1, dtype=np.uint32) * T
x = torch.from_numpy(synth.make_features(ms, 64000, seed=2)).repeat((segs * 4 * T + 63999) // 64000, 1)
x = x[:segs * 4 * T].contiguous().pin_memory().numpy()
eng.create_accumulators(tables.mean_index(ms), ms.n_means)
for _ in range(2):
    eng.score_baum_welch_host(fo, x, au, accumulate=True)
eng.kernel_timing(True)
eng.kernel_time()
t0 = time.perf_counter()
for _ in range(reps):
    eng.score_baum_welch_host(fo, x, au, accumulate=True)
eng.sync()
wall = (time.perf_counter() - t0) / reps
kms, k = eng.kernel_time()
print(f"Baum-Welch training pass: {segs * 4} x {T} frames: call {wall * 1e3:.2f} ms = {segs * 4 * T / wall / 1e6:.1f} M frames/s; "
      f"timed kernels {kms / reps:.2f} ms per call ({k // reps} launches)")
