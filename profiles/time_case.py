"""Kernel-only timing of one configuration: python profiles/time_case.py <cfg1|cfg2|cfg3> <max|sum|quant|quantscore|maxbest> T
[steps] [ld]  (ld: leading dimension of the output arrays, default M; 0 = M rounded up to a multiple of 4)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rasr_b200 import capi, synth, tables

cfg, mode, T = sys.argv[1], sys.argv[2], int(sys.argv[3])
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
ms, x = synth.make_config(cfg, T=T)
ld = int(sys.argv[5]) if len(sys.argv) > 5 else ms.M
ld = (ms.M + 3) // 4 * 4 if ld == 0 else ld
eng = capi.Engine(0)
dx = torch.from_numpy(x).cuda()
if mode in ("quant", "quantscore"):
    t = tables.quant_tables(ms)
    eng.upload_quant(t["offsets"], t["qmean"], t["cov_idx"], t["qconst"], t["isr_scaled"], t["scaling_squared"])
    q = torch.empty((T, ld), dtype=torch.int32, device="cuda")
    b = torch.empty((T, ld), dtype=torch.int32, device="cuda")
    sc = torch.empty((T, ld), dtype=torch.float32, device="cuda")
    fn = (lambda: eng.quant_score_device(T, dx, q, b, ld=ld)) if mode == "quant" else \
         (lambda: eng.quant_score_device(T, dx, None, None, sc, ld=ld))
    kern = {1: "dp4a", 2: "tensor-i8"}[eng.quant_kernel]
else:
    t = tables.float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"],
                     mode=capi.MODE_SUM if mode == "sum" else capi.MODE_MAX, kernel=capi.KERNEL_AUTO)
    s = torch.empty((T, ld), dtype=torch.float32, device="cuda")
    b = torch.empty((T, ld), dtype=torch.int32, device="cuda") if mode == "maxbest" else None
    fn = lambda: eng.score_device(T, dx, s, b, ld=ld)
    kern = {1: "direct", 2: "tensor"}[eng.float_kernel]
for _ in range(3):
    fn()
eng.sync()
eng.kernel_timing(True)
eng.kernel_time()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(steps):
    fn()
e1.record()
torch.cuda.synchronize()
kms, n = eng.kernel_time()
print(f"{cfg} {mode} T={T} ld={ld} kernel={kern}: scoring kernel {kms / steps:.4f} ms/call "
      f"({T * steps / kms / 1e3:.2f} M frames/s kernel-only), whole call incl. preparation {e0.elapsed_time(e1) / steps:.4f} ms")
