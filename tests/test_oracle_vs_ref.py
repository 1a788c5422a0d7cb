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


def _random_alignment_case(rng, integer):
    """A random automaton (cycles, dead ends, several final states, or a left-to-right HMM) and a score block."""
    from rasr_b200.align import AlignmentAutomata, linear_hmm
    M = int(rng.integers(2, 40))
    T = int(rng.integers(0, 45))
    if rng.random() < 0.4:
        seg = linear_hmm(rng.integers(0, M, int(rng.integers(1, 20))), loop=float(rng.uniform(0, 4)),
                         forward=float(rng.uniform(0, 2)), skip=float(rng.uniform(2, 30)))
    else:
        S = int(rng.integers(1, 25))
        states = [[(int(rng.integers(0, S)), int(rng.integers(0, M)),
                    float(rng.integers(0, 3)) if integer else float(rng.uniform(0, 5)))
                   for _ in range(int(rng.integers(0, 5)))] for _ in range(S)]
        final = {int(s): (float(rng.integers(0, 2)) if integer else float(rng.uniform(0, 3)))
                 for s in rng.choice(S, size=max(1, S // 3), replace=False)}
        seg = (states, int(rng.integers(0, S)), final)
    scores = (rng.integers(0, 3, (T, M)) if integer else rng.uniform(0, 80, (T, M))).astype(np.float32)
    return AlignmentAutomata.from_arc_lists([seg]), scores


@pytest.mark.skipif(not O.ref_aligner_available(), reason="oracle/_ref/ref_aligner not built")
@pytest.mark.parametrize("seed", range(40))
def test_alignment_restatements_match_compiled_reference_aligner(oracle, seed):
    """Fresh random cases through the unmodified Search::Aligner (oracle/ref_build/ref_aligner.cc; tests/golden/aln_*.npz
    hold the same on the GPU box).  Viterbi: the same emission at every frame and the same f32 score, exact ties
    (integer scores and weights, odd seeds) included.  Baum-Welch: state posteriors to 1e-5, total to 1e-6 relative."""
    rng = np.random.default_rng(1000 + seed)
    au, scores = _random_alignment_case(rng, integer=bool(seed & 1))
    T = scores.shape[0]
    fo = np.asarray([0, T], np.uint32)
    FLT_MAX = np.finfo(np.float32).max

    r = O.run_ref_aligner(scores, au, 0, mode="viterbi")
    em, st, sc = oracle.viterbi_align(fo, scores, au)
    assert np.float32(r["score"]).view(np.uint32) == sc[0].view(np.uint32)
    if sc[0] == FLT_MAX:
        assert r["time"].size == 0 and np.all(em == 0xFFFFFFFF)
    else:
        want = np.empty(T, np.uint32)
        want[r["time"]] = r["emission"]
        assert r["time"].size == T and np.array_equal(em, want) and np.all(r["weight"] == 1.0)

    r = O.run_ref_aligner(scores, au, 0, mode="baum-welch")
    w, lists, total = oracle.baum_welch_align(fo, scores, au)
    if r["score"] == FLT_MAX:
        assert total[0] == FLT_MAX and r["time"].size == 0 and not w[0].any()
        return
    assert abs(total[0] + float(r["total"])) <= 1e-6 * max(1.0, abs(total[0]))   # totalInv = MINUS the total
    assert abs(total[0] - float(r["score"])) <= 1e-5 * max(1.0, abs(total[0]))
    col = {int(e): j for j, e in enumerate(lists[0])}
    dense = np.zeros((T, max(1, len(col))), np.float64)
    for t, e, p in zip(r["time"], r["emission"], r["weight"]):
        dense[int(t), col[int(e)]] += p
    assert np.max(np.abs(dense[:, :len(col)] - w[0]), initial=0.0) <= 1e-5
