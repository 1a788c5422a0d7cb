"""GPU: the default operand split of the float tensor kernel (fp16 hi.hi plus ONE e4m3 product for both corrections,
DESIGN.md 4.1) and its packing-time guard: taken where the predicted error allows, three fp16 products otherwise
(GMM_TC_SPLIT=fp16 opts out everywhere, GMM_TC_SPLIT=fp8! bypasses the guard for experiments).  The bar is the float
scorer's: 1e-4 relative against the oracle (GaussDiagonalMaximumFeatureScorer.cc:116-218)."""
import os

import numpy as np
import pytest
import torch

from rasr_b200 import capi, synth, tables

pytestmark = pytest.mark.gpu


def _score(ms, x, split):
    old = os.environ.get("GMM_TC_SPLIT")
    if split:
        os.environ["GMM_TC_SPLIT"] = split
    else:
        os.environ.pop("GMM_TC_SPLIT", None)
    try:
        eng = capi.Engine(0)
        t = tables.float_tables(ms)
        eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX,
                         kernel=capi.KERNEL_TENSOR)
        s, _ = eng.score_host(x, want_best=False)
        kind = eng.float_split
        eng.close()
        return s, kind
    finally:
        if old is None:
            os.environ.pop("GMM_TC_SPLIT", None)
        else:
            os.environ["GMM_TC_SPLIT"] = old


def _rel(s, ref):
    return float(np.max(np.abs(s - ref) / np.abs(ref)))  # plain relative error: these scores are all above 1


def test_split_is_the_default_and_holds_the_bar_on_cfg1(oracle):
    ms, x = synth.make_config("cfg1")
    ref, _ = oracle.score_max(ms, x)
    s2, k2 = _score(ms, x, None)
    s3, k3 = _score(ms, x, "fp16")
    assert (k2, k3) == (2, 3)
    assert _rel(s3, ref) <= 1e-5
    assert _rel(s2, ref) <= 1e-4


@pytest.mark.parametrize("D,mean_scale", [(1, 2.0), (4, 2.0), (45, 6.0), (24, 3.0)])
def test_guard_keeps_three_products(oracle, D, mean_scale):
    ms = synth.make_model(M=64, K=8, D=D, seed=5, mean_scale=mean_scale)
    x = synth.make_features(ms, 64, seed=6)
    ref, _ = oracle.score_max(ms, x)
    s, kind = _score(ms, x, None)
    assert kind == 3
    assert np.max(np.abs(s - ref) / np.maximum(np.abs(ref), 1e-3)) <= 1e-4
    _, kind = _score(ms, x, "fp8!")  # the bypass for experiments
    assert kind == 2


@pytest.mark.parametrize("D", [45, 24, 9])
def test_parity_at_the_guards_edge(oracle, D):
    """The widest mean spread (in steps of 1/16 sigma) the guard still gives to the e4m3 split, against the oracle."""
    edge = None
    for i in range(8, 80):
        ms = synth.make_model(M=400, K=16, D=D, seed=7, mean_scale=i / 16.0)
        x = synth.make_features(ms, 300, seed=8)
        s, kind = _score(ms, x, None)
        if kind != 2:
            break
        edge = (i / 16.0, ms, x, s)
    assert edge is not None and edge[0] >= 1.0, "the guard refused even compact models"
    scale, ms, x, s = edge
    ref, _ = oracle.score_max(ms, x)
    err = _rel(s, ref)
    print(f"D={D}: last mean_scale on the e4m3 split {scale}, max rel err {err:.2e}")
    assert err <= 1e-4
