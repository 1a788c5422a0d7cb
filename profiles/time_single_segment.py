"""One segment per call, as MmGpu::GpuAligner hands them over: host features -> compact scores -> alignment -> path.
usage: python profiles/time_single_segment.py   (cfg1 model 4501 x 16 x 45; segments of 300 .. 4000 frames, frames / 3 labels)"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rasr_b200 import capi, synth, tables  # noqa: E402
from rasr_b200.align import AlignmentAutomata, label_slots, linear_hmm  # noqa: E402

ms = synth.make_model(4501, 16, 45, seed=1001)
t = tables.float_tables(ms)
eng = capi.Engine(0)
eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX, kernel=capi.KERNEL_TENSOR)
rng = np.random.default_rng(1)
tie = np.arange(ms.M, dtype=np.uint32)
for T in (300, 1000, 2000, 4000):
    x = synth.make_features(ms, T, seed=T)
    au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, ms.M, T // 3), loop=3.0, forward=0.0, skip=30.0)])
    slots, lo, ids, mix = label_slots(au, tie)
    for mode in ("viterbi", "baum-welch"):
        fn = (lambda: eng.score_align_labels_host([0, T], x, slots, lo, mix, want_states=False)) if mode == "viterbi" else \
             (lambda: eng.score_baum_welch_labels_host([0, T], x, slots, lo, mix))
        for _ in range(3):
            fn()
        eng.kernel_timing(True)
        eng.kernel_time()
        best = 1e9
        for _ in range(8):
            t0 = time.perf_counter()
            fn()
            best = min(best, time.perf_counter() - t0)
        kms, n = eng.kernel_time()
        print(f"{mode:10s} T={T:5d} states={T // 3 + 1:5d}: {best * 1e3:7.3f} ms per call (host wall clock, best of 8); "
              f"kernels {kms / 8:7.3f} ms in {n // 8} launches per call")
eng.close()
