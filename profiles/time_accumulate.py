"""Timing of the E-step kernel on resident frames (cfg1 model: 4501 x 16 x 45, pooled covariance).
usage: python profiles/time_accumulate.py [frames]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rasr_b200 import capi, synth, tables  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 512000
for cov in ("pooled", "density"):
    ms = synth.make_model(4501, 16, 45, seed=1001, covariance=cov)
    t = tables.float_tables(ms)
    eng = capi.Engine(0)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"])
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    x = torch.randn((T, ms.D), device="cuda")
    mix = torch.randint(0, ms.M, (T,), device="cuda", dtype=torch.int32)
    eng.accumulate_device(T, x, mix)
    eng.kernel_timing(True)
    eng.kernel_time()
    for _ in range(5):
        eng.accumulate_device(T, x, mix)
    ms_, k = eng.kernel_time()
    print(f"{cov}: {T} frames: {ms_ / k:.3f} ms per launch, {T / (ms_ / k) / 1e3:.1f} M frames/s")
    eng.close()
