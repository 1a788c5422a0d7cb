"""CPU: the oracle (oracle/gmm_oracle.c) against the golden vectors produced by the compiled reference."""
import numpy as np
import pytest

from conftest import batchint_golden_names, golden_names, load_golden


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("name", golden_names())
def test_max_scorer_bit_exact(oracle, name):
    ms, z = load_golden(name)
    s, b = oracle.score_max(ms, z["x"])
    assert np.array_equal(bits(s), bits(z["max_score"]))
    assert np.array_equal(b, z["max_best"])


@pytest.mark.parametrize("name", golden_names())
def test_sum_scorer(oracle, name):
    ms, z = load_golden(name)
    s, b = oracle.score_sum(ms, z["x"])
    # the reference's -ffast-math expf/logf differ from libm in the last bits: tolerance 1e-5 relative
    np.testing.assert_allclose(s, z["sum_score"], rtol=1e-5, atol=1e-5)
    assert np.array_equal(b, z["sum_best"])


@pytest.mark.parametrize("name", golden_names())
def test_simd_scorer_bit_exact(oracle, name):
    ms, z = load_golden(name)
    t = oracle.simd_prepare(ms)
    assert np.array_equal(t["qmean"], z["simd_qmean"])
    assert np.array_equal(t["qconst"], z["simd_qconst"])
    assert np.array_equal(t["qcov"], z["simd_qcov"])
    assert np.array_equal(bits(t["isrs"]), bits(z["simd_isrs"]))
    assert bits(np.float32(t["scaling_squared"]).reshape(1))[0] == bits(np.float32(z["simd_scaling_squared"]).reshape(1))[0]
    qx = oracle.simd_quantize_features(z["x"], t["isrs"])
    assert np.array_equal(qx, z["simd_qx"])
    q, b = oracle.simd_score(ms.mix_offsets, t, qx)
    assert np.array_equal(q, z["simd_q"])
    assert np.array_equal(b, z["simd_best"])
    assert np.array_equal(bits(oracle.simd_final_score(q, t["scaling_squared"])), bits(z["simd_score"]))


@pytest.mark.parametrize("name", batchint_golden_names())
def test_batch_int_scorer_bit_exact(oracle, name):
    """`batch-diagonal-maximum-int` (BatchFeatureScorer.cc:300-472): restated tables + the SIMD scorer's u8 distance
    + the f32 division by 2 scale^2, against the compiled reference's buffered scorer."""
    ms, z = load_golden(name)
    t = oracle.batchint_prepare(ms)
    qx = oracle.simd_quantize_features(z["x"], t["isrs"])
    q, _ = oracle.simd_score(ms.mix_offsets, t, qx)
    assert np.array_equal(bits(oracle.batchint_final_score(q, t["scale2"])), bits(z["score"]))


def test_batch_int_scorer_refuses_unpooled_covariances(oracle):
    ms, _ = load_golden("density_cov_ragged_shared")
    with pytest.raises(ValueError, match="pooled"):
        oracle.batchint_prepare(ms)
    from rasr_b200 import tables
    with pytest.raises(ValueError, match="pooled"):
        tables.batchint_tables(ms)


def test_batch_int_tables_of_the_package_match_the_oracle(oracle):
    """rasr_b200.tables.batchint_tables (numpy, exact inverse square roots) against the oracle's restatement built
    with exact inverse square roots (the reference build's rsqrt differs per build)."""
    from rasr_b200 import tables
    for name in batchint_golden_names():
        ms, _ = load_golden(name)
        t, m = oracle.batchint_prepare(ms, exact_rsqrt=True), tables.batchint_tables(ms)
        assert np.array_equal(t["qmean"], m["qmean"]) and np.array_equal(t["qconst"], m["qconst"])
        assert np.array_equal(bits(t["isrs"][0]), bits(m["isr_scaled"][0]))
        assert np.float32(m["scaling_squared"]) * np.float32(2) == t["scale2"]


def test_posteriors_sum_to_one(oracle):
    ms, z = load_golden("pooled_k16_d45")
    s, b, post = oracle.score_sum(ms, z["x"], posterior=True)
    for m in range(ms.M):
        a, e = ms.mix_offsets[m], ms.mix_offsets[m + 1]
        np.testing.assert_allclose(post[:, a:e].sum(axis=1), 1.0, rtol=1e-5)
