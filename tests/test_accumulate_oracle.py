"""CPU: the oracle's E-step restatement (oracle/gmm_oracle.c orc_accumulate) against an independent numpy
accumulation driven by the pinned scorer functions (best density of orc_score_max / posteriors of orc_score_sum)."""
import numpy as np
import pytest

from rasr_b200 import synth

SKIP = 0xFFFFFFFF


def numpy_estep(ms, x, mixture, slot_weights):
    """slot_weights: per frame a list of (slot, weight)."""
    out = dict(mean_sum=np.zeros((ms.n_means, ms.D)), mean_weight=np.zeros(ms.n_means),
               cov_sum=np.zeros((ms.n_cov, ms.D)), cov_weight=np.zeros(ms.n_cov), mix_weight=np.zeros(ms.G))
    for t, pairs in enumerate(slot_weights):
        xt = x[t].astype(np.float64)
        for g, w in pairs:
            dns = ms.mix_density[g]
            out["mean_sum"][ms.dens_mean[dns]] += w * xt
            out["cov_sum"][ms.dens_cov[dns]] += w * xt * xt
            out["mean_weight"][ms.dens_mean[dns]] += w
            out["cov_weight"][ms.dens_cov[dns]] += w
            out["mix_weight"][g] += w
    return out


@pytest.mark.parametrize("covariance", ["pooled", "density"])
@pytest.mark.parametrize("weighted", [False, True])
def test_viterbi_accumulation(oracle, covariance, weighted):
    ms = synth.make_model(M=23, K=6, D=13, seed=4, covariance=covariance)
    T = 400
    x = synth.make_features(ms, T, seed=2)
    rng = np.random.default_rng(1)
    mixture = rng.integers(0, ms.M, T).astype(np.uint32)
    mixture[rng.random(T) < 0.1] = SKIP
    weight = rng.uniform(0.1, 2.0, T) if weighted else None
    got = oracle.accumulate(ms, x, mixture, weight)
    _, best = oracle.score_max(ms, x)
    pairs = [[] if m == SKIP else [(int(ms.mix_offsets[m] + best[t, m]), 1.0 if weight is None else weight[t])]
             for t, m in enumerate(mixture)]
    want = numpy_estep(ms, x, mixture, pairs)
    keep = mixture != SKIP
    assert np.array_equal(got["best"][keep], best[np.arange(T)[keep], mixture[keep]])
    assert np.all(got["best"][~keep] == SKIP)
    for k in want:
        np.testing.assert_allclose(got[k], want[k], rtol=1e-13, atol=0)
    if not weighted:
        assert got["mix_weight"].sum() == keep.sum() and np.array_equal(got["mix_weight"], want["mix_weight"])


def test_posterior_weighted_accumulation(oracle):
    ms = synth.make_model(M=11, K=5, D=9, seed=8, covariance="mixture")
    T = 200
    x = synth.make_features(ms, T, seed=3)
    rng = np.random.default_rng(2)
    mixture = rng.integers(0, ms.M, T).astype(np.uint32)
    weight = rng.uniform(0.5, 1.5, T)
    thr = float(np.finfo(np.float32).eps)
    got = oracle.accumulate(ms, x, mixture, weight, sum_scorer=True, viterbi=False, weight_threshold=thr)
    _, best, post = oracle.score_sum(ms, x, posterior=True)
    pairs = []
    for t, m in enumerate(mixture):
        a, b = int(ms.mix_offsets[m]), int(ms.mix_offsets[m + 1])
        pairs.append([(g, weight[t] * float(post[t, g])) for g in range(a, b) if weight[t] * float(post[t, g]) > thr])
    want = numpy_estep(ms, x, mixture, pairs)
    assert np.array_equal(got["best"], best[np.arange(T), mixture])
    for k in want:
        np.testing.assert_allclose(got[k], want[k], rtol=1e-13, atol=0)
    # posteriors of a mixture sum to one: total weight = sum of frame weights (up to f32 exp rounding)
    np.testing.assert_allclose(got["mix_weight"].sum(), weight.sum(), rtol=1e-5)
