"""CPU emulation of the float tensor kernel's operand splits on the cfg1 model (4501 x 16 x 45, 200 frames): how far
the per-mixture minimum moves when the whitened operands go to the tensor core as fp16 parts (products accumulated
exactly here; the kernel accumulates in f32).  Run: python profiles/split_error.py
Result (this round): three products 2.2e-07, two products 4.6e-04..4.8e-04 (0.5 % of the scores beyond the 1e-4 bar),
one product 7.4e-04 -- the three-product split is the cheapest one that holds the bar."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from rasr_b200 import synth, tables

ms, x = synth.make_config("cfg1")
t = tables.float_tables(ms)
mean, isr, w2, ln, off = t["mean"], t["isr"], t["w2"], t["log_norm"], t["offsets"]
x = x[:200].astype(np.float32)
c = mean.mean(0)
xt = ((x - c) * isr[0]).astype(np.float32)     # whitened frames (pooled covariance)
mt = ((mean - c) * isr[0]).astype(np.float32)  # whitened means
const = (mt.astype(np.float64) ** 2).sum(1) + w2.astype(np.float64) + ln[t["cov_idx"]].astype(np.float64)
x2 = (xt.astype(np.float64) ** 2).sum(1)[:, None]


def split(a):
    hi = a.astype(np.float16)
    lo = (a - hi.astype(np.float32)).astype(np.float16)
    return hi.astype(np.float64), lo.astype(np.float64)


def mixture_min(dot):
    s = 0.5 * (x2 - 2 * dot + const[None, :])
    return np.stack([s[:, off[m]:off[m + 1]].min(1) for m in range(len(off) - 1)], 1)


xh, xl = split(xt)
mh, ml = split(mt)
ref = mixture_min(xt.astype(np.float64) @ mt.astype(np.float64).T)
for name, dot in (("three products (hi.hi + lo.hi + hi.lo)", xh @ mh.T + xl @ mh.T + xh @ ml.T),
                  ("two products, model in one part", xh @ mh.T + xl @ mh.T),
                  ("two products, frames in one part", xh @ mh.T + xh @ ml.T),
                  ("one product", xh @ mh.T)):
    rel = np.abs(mixture_min(dot) - ref) / np.maximum(np.abs(ref), 1)
    print(f"{name:40s} max rel err {rel.max():.3e}   mean {rel.mean():.3e}   beyond 1e-4: {(rel > 1e-4).mean():.4f}")
