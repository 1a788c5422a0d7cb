"""How the e5m2-correction split (GMM_TC_SPLIT=fp8) degrades with the spread of the means: emulation as in
split_error.py on 400 x 16 x D models, frames drawn near the densities.  kappa/D = largest squared whitened norm of
a centred mean per dimension (the packing-time quantity the 3-product split guards at 64).
Run: python profiles/split_error_sweep.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from rasr_b200 import synth, tables


def e5m2(a):
    a = np.asarray(a, np.float64)
    m, e = np.frexp(a)
    q = np.ldexp(np.round(m * 8) / 8, e)
    q = np.where(np.abs(a) < 2.0 ** -14, np.round(a * 2.0 ** 16) / 2.0 ** 16, q)
    return np.clip(q, -57344.0, 57344.0)


def split(a):
    hi = a.astype(np.float16)
    return hi.astype(np.float64), (a - hi.astype(np.float32)).astype(np.float64)


print(f"{'D':>3} {'mean_scale':>10} {'kappa/D':>8} {'3 products':>11} {'fp16 + e5m2':>12} {'beyond 1e-4':>12}")
for D in (1, 4, 9, 24, 45):
    for scale in (0.5, 1.0, 2.0, 3.0, 4.0, 6.0):
        ms = synth.make_model(M=400, K=16, D=D, seed=7, mean_scale=scale)
        x = synth.make_features(ms, 200, seed=8).astype(np.float32)
        t = tables.float_tables(ms)
        mean, isr, w2, ln, off = t["mean"], t["isr"], t["w2"], t["log_norm"], t["offsets"]
        c = mean.mean(0)
        xt = ((x - c) * isr[0]).astype(np.float32)
        mt = ((mean - c) * isr[0]).astype(np.float32)
        const = (mt.astype(np.float64) ** 2).sum(1) + w2.astype(np.float64) + ln[t["cov_idx"]].astype(np.float64)
        x2 = (xt.astype(np.float64) ** 2).sum(1)[:, None]

        def mixture_min(dot):
            s = 0.5 * (x2 - 2 * dot + const[None, :])
            return np.stack([s[:, off[m]:off[m + 1]].min(1) for m in range(len(off) - 1)], 1)

        ref = mixture_min(xt.astype(np.float64) @ mt.astype(np.float64).T)
        (xh, xl), (mh, ml) = split(xt), split(mt)
        xl16, ml16 = xl.astype(np.float16).astype(np.float64), ml.astype(np.float16).astype(np.float64)
        r3 = np.abs(mixture_min(xh @ mh.T + xl16 @ mh.T + xh @ ml16.T) - ref) / np.maximum(np.abs(ref), 1)
        r8 = np.abs(mixture_min(xh @ mh.T + e5m2(xl) @ e5m2(mh).T + e5m2(xh) @ e5m2(ml).T) - ref) / np.maximum(np.abs(ref), 1)
        kap = (mt.astype(np.float64) ** 2).sum(1).max() / D
        print(f"{D:3d} {scale:10.1f} {kap:8.1f} {r3.max():11.2e} {r8.max():12.2e} {(r8 > 1e-4).mean():12.4f}")
