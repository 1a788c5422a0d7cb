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


@pytest.mark.skipif(not O.ref_estimator_available(), reason="oracle/_ref/ref_estimator not built")
@pytest.mark.parametrize("kw", [dict(M=23, K=8, D=33, seed=31), dict(M=9, K=4, D=12, covariance="mixture", seed=32)])
def test_estimator_restatements_match_compiled_reference(tmp_path, oracle, kw):
    """Fresh random case through the unmodified estimator stack (the golden est_*.npz hold the same on the GPU box):
    accumulator files bit for bit in viterbi mode, the text model writer byte for byte, estimate() as printed."""
    from rasr_b200 import model_io
    ms = synth.make_model(**kw)
    T = 900
    x = synth.make_features(ms, T, seed=7)
    rng = np.random.default_rng(kw["seed"])
    mixture = np.repeat(rng.integers(0, ms.M, T // 3), 3).astype(np.uint32)
    mixture[: ms.M * 6] = np.repeat(np.arange(ms.M), 6)
    weight = rng.uniform(0.1, 3.0, T)
    for scorer, weighted in (("max", False), ("max", True), ("sum", True)):
        r = O.run_ref_estimator(ms, x, mixture, weight if weighted else None, mode="viterbi", scorer=scorer)
        acc = oracle.accumulate(ms, x, mixture, weight if weighted else None, sum_scorer=scorer == "sum")
        est = model_io.MixtureSetEstimatorData.from_accumulators(ms, acc)
        model_io.write_mixset(est, tmp_path / "a.mixset")
        assert (tmp_path / "a.mixset").read_bytes() == r["mixset"]
        assert np.array_equal(acc["best"], r["best"])
        model_io.write_mixture_set_text(est.estimate(), tmp_path / "e.pms")
        assert (tmp_path / "e.pms").read_bytes() == r["est_pms"]
    model_io.write_mixture_set_text(ms, tmp_path / "m.pms")
    assert (tmp_path / "m.pms").read_bytes() == r["pms"] == r["reread_pms"]
