"""CPU, build container only: the oracle against the compiled reference on fresh random models (skipped when
oracle/_ref/ref_scorer is absent).  The golden fixtures cover the same comparison on the GPU box."""
import numpy as np
import pytest

from oracle import oracle as O
from rasr_b200 import synth

pytestmark = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref/ref_scorer not built")

CASES = [dict(M=37, K=16, D=45), dict(M=9, K=5, D=17, covariance="density", ragged=True, shared=True,
                                      zero_weight_frac=0.2), dict(M=11, K=3, D=4, covariance="mixture"),
         dict(M=3, K=64, D=45, seed=5)]


@pytest.mark.parametrize("kw", CASES)
def test_oracle_matches_compiled_reference(oracle, kw):
    ms = synth.make_model(**kw)
    x = synth.make_features(ms, 40)
    r = O.run_ref(ms, x, "max")
    s, b = oracle.score_max(ms, x)
    assert np.array_equal(r["score"].view(np.uint32), s.view(np.uint32)) and np.array_equal(r["best"], b)
    r = O.run_ref(ms, x, "sum")
    s, b = oracle.score_sum(ms, x)
    np.testing.assert_allclose(s, r["score"], rtol=1e-5, atol=1e-5)
    r = O.run_ref(ms, x, "simd")
    t = oracle.simd_prepare(ms)
    qx = oracle.simd_quantize_features(x, t["isrs"])
    q, b = oracle.simd_score(ms.mix_offsets, t, qx)
    assert np.array_equal(r["q"], q) and np.array_equal(r["best"], b) and np.array_equal(r["qx"], qx)
    assert np.array_equal(r["qmean"], t["qmean"]) and np.array_equal(r["qconst"], t["qconst"])
    assert np.array_equal(r["score"].view(np.uint32), oracle.simd_final_score(q, t["scaling_squared"]).view(np.uint32))
