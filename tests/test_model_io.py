"""CPU: RASR text mixture sets, MIXSET estimator files and the M-step (rasr_b200/model_io.py)."""
import gzip
import struct

import numpy as np
import pytest

from rasr_b200 import model_io, synth, tables
from rasr_b200.gms import MixtureSetData

SAMPLE = """#Version: 2.0
#CovarianceType: DiagonalCovariance
2 3 4 3 2
2 0 -0.693147 1 -0.693147
1 2 0
0
0 0
1 0
2 1
2 1
2 1.5 -2
2 0.25 3e+06
2 -1 0
 2 1 1 4 1
 2 0.5 2 2 0.25
"""


def assert_same_model(a: MixtureSetData, b: MixtureSetData):
    for f in ("mix_offsets", "mix_density", "mix_logw", "dens_mean", "dens_cov", "means", "variances"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f
    assert a.D == b.D


def test_read_text_as_the_reference_writes_it(tmp_path):
    p = tmp_path / "am.pms"
    p.write_text(SAMPLE)
    ms = model_io.read_mixture_set_text(p)
    assert (ms.D, ms.M, ms.n_densities, ms.n_means, ms.n_cov) == (2, 3, 4, 3, 2)
    assert list(ms.mix_offsets) == [0, 2, 3, 3]                     # third mixture: no densities
    assert list(ms.mix_density) == [0, 1, 2] and ms.mix_logw[0] == -0.693147 and ms.mix_logw[2] == 0
    assert list(ms.dens_mean) == [0, 1, 2, 2] and list(ms.dens_cov) == [0, 0, 1, 1]
    assert np.array_equal(ms.means, np.array([[1.5, -2], [0.25, 3e6], [-1, 0]], np.float32))
    # variance = value * feature weight (DiagonalCovariance::read)
    assert np.array_equal(ms.variances, np.array([[1, 4], [1.0, 0.5]], np.float32))
    gz = tmp_path / "am.pms.gz"
    with gzip.open(gz, "wt") as f:
        f.write(SAMPLE)
    assert_same_model(ms, model_io.read_mixture_set_text(gz))


def test_text_version_1_stores_weights(tmp_path):
    p = tmp_path / "old.pms"
    p.write_text(SAMPLE.replace("#Version: 2.0", "#Version: 1.0").replace("-0.693147", "0.5").replace("1 2 0\n", "1 2 0\n", 1))
    ms = model_io.read_mixture_set_text(p)
    assert ms.mix_logw[0] == np.log(0.5) and ms.mix_logw[2] == model_io.F64_MIN      # weight 0 -> f64 min
    p.write_text(SAMPLE.replace("2.0", "3.0"))
    with pytest.raises(ValueError, match="not supported"):
        model_io.read_mixture_set_text(p)
    p.write_text(SAMPLE.replace("DiagonalCovariance", "FullCovariance"))
    with pytest.raises(ValueError, match="Covariance Type"):
        model_io.read_mixture_set_text(p)


@pytest.mark.parametrize("covariance", ["pooled", "density", "mixture"])
def test_text_round_trip(tmp_path, covariance):
    ms = synth.make_model(M=19, K=4, D=7, seed=3, covariance=covariance)
    p = tmp_path / "m.pms.gz"
    model_io.write_mixture_set_text_exact(ms, p)
    assert_same_model(ms, model_io.read_mixture_set_text(p))
    # the reference's own writer keeps 6 significant digits
    q = tmp_path / "m6.pms"
    model_io.write_mixture_set_text(ms, q)
    back = model_io.read_mixture_set_text(q)
    np.testing.assert_allclose(back.means, ms.means, rtol=1e-5, atol=1e-30)
    np.testing.assert_allclose(back.variances, ms.variances, rtol=1e-5)
    assert np.array_equal(back.mix_density, ms.mix_density)
    # and the GPU tables built from the file are the tables built from the model
    t1, t2 = tables.float_tables(ms), tables.float_tables(model_io.read_mixture_set_text(p))
    assert all(np.array_equal(t1[k], t2[k]) for k in t1)


def estimator_for(ms, oracle, T=4000, seed=0, weighted=False):
    x = synth.make_features(ms, T, seed=seed)
    rng = np.random.default_rng(seed)
    mixture = rng.integers(0, ms.M, T).astype(np.uint32)
    w = rng.uniform(0.5, 2, T) if weighted else None
    acc = oracle.accumulate(ms, x, mixture, w)
    return model_io.MixtureSetEstimatorData.from_accumulators(ms, acc), x, mixture, acc


def test_mixset_round_trip_and_layout(tmp_path, oracle):
    ms = synth.make_model(M=9, K=3, D=5, seed=2, covariance="mixture")
    est, *_ = estimator_for(ms, oracle, weighted=True)
    p = tmp_path / "acc.mix"
    model_io.write_mixset(est, p)
    raw = p.read_bytes()
    assert raw[:8] == b"MIXSET\0\0" and struct.unpack_from("<II", raw, 8) == (2, 5)
    n_means = struct.unpack_from("<I", raw, 16)[0]
    assert n_means == ms.n_means and struct.unpack_from("<I", raw, 20)[0] == 5     # first accumulator: size D
    back = model_io.read_mixset(p)
    assert back.equal_topology(est)                     # synth numbers estimators in first-appearance order
    for f in ("mix_weight", "mean_sum", "mean_weight", "cov_sum", "cov_weight"):
        assert np.array_equal(getattr(back, f), getattr(est, f)), f
    bad = tmp_path / "bad.mix"
    bad.write_bytes(b"MIXSETX\0" + raw[8:])
    with pytest.raises(ValueError, match="magic"):
        model_io.read_mixset(bad)


def test_mixset_version_0_counts(tmp_path):
    D = 2
    b = b"MIXSET\0\0" + struct.pack("<II", 0, D)
    b += struct.pack("<I", 1) + struct.pack("<I", D) + struct.pack("<2d", 3.0, -1.0) + struct.pack("<I", 4)
    b += struct.pack("<I", 1) + struct.pack("<I", D) + struct.pack("<2d", 5.0, 2.0) + struct.pack("<I", 4)
    b += struct.pack("<I", 1) + struct.pack("<II", 0, 0)
    b += struct.pack("<I", 1) + struct.pack("<I", 1) + struct.pack("<II", 0, 4)
    p = tmp_path / "v0.mix"
    p.write_bytes(b)
    est = model_io.read_mixset(p)
    assert est.mean_weight[0] == 4 and est.mix_weight[0] == 4 and list(est.mean_sum[0]) == [3, -1]
    ms = est.estimate(min_observation_weight=0)
    assert np.array_equal(ms.means, np.array([[0.75, -0.25]], np.float32))
    assert np.allclose(ms.variances, [[5 / 4 - 0.75 ** 2, 2 / 4 - 0.25 ** 2]])
    assert ms.mix_logw[0] == 0


def test_combine_and_m_step_against_numpy(oracle):
    ms = synth.make_model(M=12, K=4, D=6, seed=5, covariance="pooled")
    e1, x1, m1, _ = estimator_for(ms, oracle, T=3000, seed=1)
    e2, x2, m2, _ = estimator_for(ms, oracle, T=2000, seed=2)
    assert e1.accumulate(e2)
    x, mix = np.concatenate([x1, x2]).astype(np.float64), np.concatenate([m1, m2])
    _, best = oracle.score_max(ms, np.concatenate([x1, x2]))
    slot = ms.mix_offsets[mix] + best[np.arange(len(mix)), mix]
    new = e1.estimate(min_observation_weight=0)
    assert new.M == ms.M and new.n_cov == 1
    # mixture weights: relative counts; means: averages of the assigned frames; pooled variance: within-class
    counts = np.bincount(slot, minlength=ms.G).astype(np.float64)
    total, within = 0.0, np.zeros(ms.D)
    for m in range(ms.M):
        a, b = int(ms.mix_offsets[m]), int(ms.mix_offsets[m + 1])
        c = counts[a:b]
        na, nb = int(new.mix_offsets[m]), int(new.mix_offsets[m + 1])
        assert nb - na == b - a                                      # nothing removed at min weight 0
        with np.errstate(divide="ignore"):
            want = np.where(c > 0, np.log(np.maximum(c, 1e-300) / c.sum()), -np.inf)
        got = new.mix_logw[na:nb]
        assert np.allclose(got[c > 0], want[c > 0], rtol=0, atol=1e-12) and np.all(got[c == 0] < -1e300)
        for k in range(b - a):
            sel = slot == a + k
            if sel.any():
                mu = x[sel].mean(axis=0)
                mean_row = new.means[new.dens_mean[new.mix_density[na + k]]]
                np.testing.assert_allclose(mean_row, mu, rtol=2e-6, atol=1e-6)
                within += ((x[sel] - mu) ** 2).sum(axis=0)
                total += sel.sum()
    np.testing.assert_allclose(new.variances[0], within / total, rtol=1e-5)
    other = synth.make_model(M=12, K=3, D=6, seed=5, covariance="pooled")
    e3, *_ = estimator_for(other, oracle, T=100)
    assert not e1.accumulate(e3)


def test_m_step_removes_light_densities_but_keeps_the_heaviest(oracle):
    ms = synth.make_model(M=6, K=5, D=4, seed=7, covariance="density")
    est, *_ = estimator_for(ms, oracle, T=300)
    new = est.estimate()                                             # reference default: minimum observation weight 5
    for m in range(ms.M):
        a, b = int(ms.mix_offsets[m]), int(ms.mix_offsets[m + 1])
        w = est.mix_weight[a:b]
        keep = (w >= 5) | (np.arange(b - a) == int(np.argmax(w)))
        assert int(new.mix_offsets[m + 1] - new.mix_offsets[m]) == int(keep.sum())
        lw = new.mix_logw[new.mix_offsets[m]:new.mix_offsets[m + 1]]
        assert abs(np.exp(lw).sum() - 1) < 1e-12
    assert new.n_densities == int(new.mix_offsets[-1]) <= ms.n_densities
    assert new.n_means == new.n_densities and new.n_cov == new.n_densities
    est.mix_weight[:] = 0
    with pytest.raises(ValueError, match="zero weight"):
        est.estimate()
    est.estimate(allow_zero_weights=True)


def test_m_step_may_return_zero_variances_like_the_reference(tmp_path):
    """minimum-variance defaults to 0 (AbstractMixtureSetEstimator.cc:35-38): a covariance estimated from identical
    observations comes out as 0, the set can still be written and read; only building a scorer from it is refused
    (CovarianceFeatureScorerElement.cc:23,41-44)."""
    est = model_io.MixtureSetEstimatorData(2, [0, 1], [0], [6.0], [0], [0], [[6.0, 12.0]], [6.0], [[6.0, 24.0]], [6.0])
    ms = est.estimate()
    assert np.array_equal(ms.variances, np.zeros((1, 2), np.float32))
    model_io.write_mixture_set_text(ms, tmp_path / "z.pms")
    assert np.array_equal(model_io.read_mixture_set_text(tmp_path / "z.pms").variances, ms.variances)
    with pytest.raises(ValueError, match="strictly positive"):
        tables.float_tables(ms)
    assert np.all(est.estimate(min_variance=0.01).variances == np.float32(0.01))
