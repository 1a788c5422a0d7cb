"""Timing of the front-end kernel (window 9 x 16 -> 45) on resident base features.
usage: python profiles/time_lda.py [segments] [frames_per_segment]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rasr_b200 import capi  # noqa: E402

segs = int(sys.argv[1]) if len(sys.argv) > 1 else 512
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
eng = capi.Engine(0)
rng = np.random.default_rng(0)
eng.set_lda(rng.standard_normal((45, 144)).astype(np.float32), 16, 9, 4)
fo = np.arange(segs + 1, dtype=np.uint32) * T
base = torch.randn((segs * T, 16), device="cuda")
out = torch.empty((segs * T, 45), device="cuda")
eng.lda_device(fo, base, out)
eng.kernel_timing(True)
eng.kernel_time()
for _ in range(5):
    eng.lda_device(fo, base, out)
ms, k = eng.kernel_time()
n = segs * T
print(f"lda 45x(16*9): {n} frames: {ms / k:.3f} ms per launch, {n / (ms / k) / 1e3:.1f} M frames/s, "
      f"{n * (16 + 45) * 4 / (ms / k) / 1e6:.1f} GB/s algorithmic")
