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


# ---- what the correction products need: hi.hi in fp16, lo.hi and hi.lo with BOTH operands rounded to e4m3's four
# significant bits (range handled by a power-of-two block scale, as kind::mxf8f6f4 provides) -- an 8-bit MMA runs at
# twice the fp16 rate, so this split would cost 2 fp16-product times instead of 3
def sig_bits(a, bits):
    m, e = np.frexp(a)
    return np.ldexp(np.round(m * (1 << bits)) / (1 << bits), e)


for bits, label in ((4, "e4m3 (4 significant bits)"), (3, "e5m2 (3 significant bits)")):
    dot = xh @ mh.T + sig_bits(xl, bits) @ sig_bits(mh, bits).T + sig_bits(xh, bits) @ sig_bits(ml, bits).T
    rel = np.abs(mixture_min(dot) - ref) / np.maximum(np.abs(ref), 1)
    print(f"{'fp16 hi.hi + corrections in ' + label:60s} max rel err {rel.max():.3e}   beyond 1e-4: {(rel > 1e-4).mean():.4f}")


# the same with e5m2's RANGE as well (no block scale: plain kind::f8f6f4): normal numbers from 2^-14, subnormal step
# 2^-16, round to nearest; the lo parts of small operands underflow, which only matters in absolute terms
def e5m2(a):
    a = np.asarray(a, np.float64)
    m, e = np.frexp(a)                                   # a = m * 2^e, 0.5 <= |m| < 1
    q = np.ldexp(np.round(m * 8) / 8, e)                 # 3 significant bits
    sub = np.abs(a) < 2.0 ** -14
    q = np.where(sub, np.round(a * 2.0 ** 16) / 2.0 ** 16, q)
    return np.clip(q, -57344.0, 57344.0)


dot = xh @ mh.T + e5m2(xl) @ e5m2(mh).T + e5m2(xh) @ e5m2(ml).T
rel = np.abs(mixture_min(dot) - ref) / np.maximum(np.abs(ref), 1)
print(f"{'fp16 hi.hi + corrections in e5m2 incl. its range':60s} max rel err {rel.max():.3e}   beyond 1e-4: {(rel > 1e-4).mean():.4f}")
print("smallest / largest |lo| of the frames:", np.abs(xl[xl != 0]).min(), np.abs(xl).max(), " of the means:", np.abs(ml[ml != 0]).min(), np.abs(ml).max())
