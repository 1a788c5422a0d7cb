"""Small driver for ncu captures: python profiles/prof_case.py <cfg1|cfg2|cfg3> <max|sum|quant> [T] [launches]."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rasr_b200 import capi, synth, tables

cfg, mode = sys.argv[1], sys.argv[2]
T = int(sys.argv[3]) if len(sys.argv) > 3 else None
n = int(sys.argv[4]) if len(sys.argv) > 4 else 3
ms, x = synth.make_config(cfg, T=T)
T = x.shape[0]
eng = capi.Engine(0)
dx = torch.from_numpy(x).cuda()
if mode == "quant":
    t = tables.quant_tables(ms)
    eng.upload_quant(t["offsets"], t["qmean"], t["cov_idx"], t["qconst"], t["isr_scaled"], t["scaling_squared"])
    ld = (ms.M + 3) // 4 * 4  # rows padded to 16 bytes, as bench.py
    q = torch.empty((T, ld), dtype=torch.int32, device="cuda")
    b = torch.empty((T, ld), dtype=torch.int32, device="cuda")
    for _ in range(n):
        eng.quant_score_device(T, dx, q, b, ld=ld)
else:
    t = tables.float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"],
                     mode=capi.MODE_SUM if mode == "sum" else capi.MODE_MAX, kernel=capi.KERNEL_AUTO)
    ld = (ms.M + 3) // 4 * 4
    s = torch.empty((T, ld), dtype=torch.float32, device="cuda")
    for _ in range(n):
        eng.score_device(T, dx, s, None, ld=ld)
eng.sync()
torch.cuda.synchronize()
print("ok", cfg, mode, T, "kernel", eng.float_kernel if mode != "quant" else "quant")
