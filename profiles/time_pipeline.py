"""Timing of the features-in / paths-out pipeline (gmm_score_align_host) with full and compact score blocks:
whole call (wall clock, pinned host features) and the timed kernels inside (scoring + alignment).
usage: python profiles/time_pipeline.py [segments] [frames_per_segment] [emissions_per_automaton]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rasr_b200 import capi, synth, tables  # noqa: E402
from rasr_b200.align import AlignmentAutomata, linear_hmm  # noqa: E402

segs = int(sys.argv[1]) if len(sys.argv) > 1 else 512
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
n = int(sys.argv[3]) if len(sys.argv) > 3 else 300
ms = synth.make_model(M=4501, K=16, D=45, seed=1001)
eng = capi.Engine(0)
t = tables.float_tables(ms)
eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX,
                 kernel=capi.KERNEL_AUTO)
rng = np.random.default_rng(0)
au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, ms.M, n), 3.0, 0.0, 30.0) for _ in range(segs)])
fo = np.arange(segs + 1, dtype=np.uint32) * T
x = torch.from_numpy(synth.make_features(ms, min(segs * T, 64000), seed=2)).repeat((segs * T + 63999) // 64000, 1)
x = x[:segs * T].contiguous().pin_memory().numpy()
eng.create_accumulators(tables.mean_index(ms), ms.n_means)
ref = None
cases = [("0", "100000"), ("1", "100000"), ("1", "296"), ("1", "148"), ("1", "74")]
if os.environ.get("PIPELINE_CASES"):  # e.g. "1:296" -- compact blocks, groups of 296 segments (the defaults) only
    cases = [tuple(c.split(":")) for c in os.environ["PIPELINE_CASES"].split(",")]
for flag, group in cases:
    os.environ["GMM_ALIGN_SUBSET"] = flag
    os.environ["GMM_ALIGN_GROUP"] = group
    for acc in (False, True):
        for _ in range(2):
            out = eng.score_align_host(fo, x, au, want_states=False, accumulate=acc, want_paths=not acc)
        eng.kernel_timing(True)
        eng.kernel_time()
        t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            out = eng.score_align_host(fo, x, au, want_states=False, accumulate=acc, want_paths=not acc)
        eng.sync()
        wall = (time.perf_counter() - t0) / reps
        kms, k = eng.kernel_time()
        eng.kernel_timing(False)
        print(f"subset={flag} group={group} accumulate={acc}: {segs} x {T} frames, {n} emissions: call {wall * 1e3:.2f} ms = "
              f"{segs * T / wall / 1e6:.1f} M frames/s; timed kernels {kms / reps:.2f} ms per call ({k // reps} launches)")
        if not acc:
            if ref is None:
                ref = out
            else:
                assert np.array_equal(ref[0], out[0]) and np.array_equal(ref[2].view(np.uint32), out[2].view(np.uint32))
                print("paths and path scores identical with full and compact score blocks")
