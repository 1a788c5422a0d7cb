"""GPU parity tests: the CUDA path, called through the C-ABI (rasr_b200.capi), against the oracle and the
golden vectors of the compiled reference.

Bars: quantized scorer bit-exact (q, argmin, f32 score, quantized frames); float direct kernel bit-exact for the
maximum approximation when fed the reference's tables, float tensor kernel and sum scorer within 1e-4 relative
on the -log scores (BASELINE.json north_star), arg-min equal wherever the runner-up is not within tolerance.
"""
import numpy as np
import pytest
import torch

from conftest import (batchint_golden_names, golden_names, load_golden, oracle_batchint_tables, oracle_float_tables,
                      oracle_quant_tables, ref_quant_tables)
from rasr_b200 import capi, feature_scorer, synth, tables

pytestmark = pytest.mark.gpu
RTOL = 1e-4  # north_star: "within 1e-4 relative on -log scores for the float scorer"


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def eng():
    e = capi.Engine(0)
    yield e
    e.close()


def upload_float(eng, t, mode=capi.MODE_MAX, kernel=capi.KERNEL_DIRECT):
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=mode, kernel=kernel)


def upload_quant(eng, t):
    eng.upload_quant(t["offsets"], t["qmean"], t["cov_idx"], t["qconst"], t["isr_scaled"], t["scaling_squared"])


def assert_float_close(s, ref, rtol=RTOL):
    """Relative error against max(|ref|, 1).  The floor is not slack for the kernel: a score is the sum of a
    log-normalisation term, a weight term and a distance, each of magnitude 10..10^3 and of either sign, so scores
    cross zero, where a relative error is undefined for the reference's own f32 evaluation as well (its rounding
    error is ~1e-7 of the TERMS, not of the sum).  Below |score| = 1 the bar therefore reads 1e-4 absolute -- still
    1e-6 of the terms -- and at or above 1 (all but a handful of values in these tests) it is the north star's
    1e-4 relative."""
    fin = np.isfinite(ref) & (np.abs(ref) < 1e37)
    assert np.array_equal(fin, np.isfinite(s) & (np.abs(s) < 1e37))
    err = np.abs(s[fin] - ref[fin]) / np.maximum(np.abs(ref[fin]), 1.0)
    assert err.size == 0 or err.max() <= rtol, f"max rel err {err.max():.3e}"
    if (~fin).any():  # mixtures without a finite density: 0.5 * FLT_MAX in the reference
        assert np.array_equal(bits(s[~fin]), bits(ref[~fin]))


def assert_best_consistent(ms, x, best, ref_best, oracle, tab):
    """arg-min must equal the reference's except where the two candidates' scores tie within tolerance."""
    diff = np.argwhere(best != ref_best)
    if diff.size == 0:
        return
    isr, logn, w2 = tab
    mean, cov = tables.flatten(ms)
    for t, m in diff[:200]:
        a = int(ms.mix_offsets[m])
        def sc(k):
            g = a + int(k)
            d = (((mean[g].astype(np.float64) - x[t]) * isr[cov[g]]) ** 2).sum()
            return w2[g] + logn[cov[g]] + d
        s1, s2 = sc(best[t, m]), sc(ref_best[t, m])
        assert abs(s1 - s2) <= 2 * RTOL * max(abs(s2), 1.0), (t, m, best[t, m], ref_best[t, m], s1, s2)


# ---------------------------------------------------------------------------------------------------
# golden vectors of the compiled reference
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", golden_names())
def test_golden_quantized_bit_exact(eng, name):
    ms, z = load_golden(name)
    upload_quant(eng, ref_quant_tables(ms, z))
    r = eng.quant_score_host(z["x"], want_qx=True)
    assert np.array_equal(r["qx"], z["simd_qx"])
    assert np.array_equal(r["q"], z["simd_q"])
    assert np.array_equal(r["best"], z["simd_best"])
    assert np.array_equal(bits(r["score"]), bits(z["simd_score"]))


@pytest.mark.parametrize("name", batchint_golden_names())
def test_golden_batch_int_bit_exact(eng, oracle, name):
    """`batch-diagonal-maximum-int` (BatchFeatureScorer.cc:300-472) on the quantized kernel: the reference's own
    constants and final scale, scores bit-identical to the compiled reference's buffered scorer -- including the
    zero-weight densities whose 0x80000000 constant wins the minimum there."""
    ms, z = load_golden(name)
    t, raw = oracle_batchint_tables(ms)
    eng.upload_quant(t["offsets"], t["qmean"], t["cov_idx"], t["qconst"], t["isr_scaled"], t["scaling_squared"],
                     rule=capi.QUANT_RULE_BATCH_INT)
    r = eng.quant_score_host(z["x"])
    assert np.array_equal(bits(r["score"]), bits(z["score"]))
    qx = oracle.simd_quantize_features(z["x"], raw["isrs"])
    q, best = oracle.simd_score(ms.mix_offsets, raw, qx)
    assert np.array_equal(r["q"], q) and np.array_equal(r["best"], best)


def test_density_posteriors_through_the_host_mirror(oracle):
    """getDensityPosteriorProbabilities of a sum scorer's context (the plugin's binding, AssigningFeatureScorer.hh:47)
    against the oracle's posteriors; the maximum scorer refuses like the reference."""
    ms = synth.make_model(M=40, K=9, D=21, seed=31, ragged=True)
    x = synth.make_features(ms, 12, seed=2)
    fs = feature_scorer.GpuFeatureScorer(ms, "diagonal-sum", float_tables=oracle_float_tables(ms))
    fs.reset()
    for f in x:
        fs.addFeature(f)
    ctx = [fs.flush() for _ in range(len(x))]
    _, _, post = oracle.score_sum(ms, x, posterior=True)
    for t in (0, 5, 11):
        for e in (0, 7, 39):
            a, b = ms.mix_offsets[e], ms.mix_offsets[e + 1]
            np.testing.assert_allclose(ctx[t].getDensityPosteriorProbabilities(e), post[t, a:b], rtol=1e-4, atol=1e-7)
    assert fs.n_batches == 1
    with pytest.raises(ValueError, match="density in the mixture"):
        ctx[0].scoreOfDensity(0, 0)
    fm = feature_scorer.GpuFeatureScorer(ms, "diagonal-maximum")
    with pytest.raises(ValueError, match="posterior"):
        fm.getScorer(x[0]).getDensityPosteriorProbabilities(0)


def test_batch_int_scorer_type_of_the_host_mirror(oracle):
    """feature-scorer-type = batch-diagonal-maximum-int through the buffered interface, at a size where the tensor
    kernel runs: equal to the oracle's restatement of the reference scorer."""
    ms = synth.make_model(M=301, K=16, D=45, seed=8)
    x = synth.make_features(ms, 200, seed=3)
    t, raw = oracle_batchint_tables(ms)
    fs = feature_scorer.GpuFeatureScorer(ms, "batch-diagonal-maximum-int", quant_tables=t)
    fs.reset()
    for f in x:
        fs.addFeature(f)
    got = np.stack([fs.flush().scores() for _ in range(len(x))])
    q, _ = oracle.simd_score(ms.mix_offsets, raw, oracle.simd_quantize_features(x, raw["isrs"]))
    assert np.array_equal(bits(got), bits(oracle.batchint_final_score(q, raw["scale2"])))
    assert fs.n_batches == 1
    with pytest.raises(ValueError, match="pooled"):
        feature_scorer.GpuFeatureScorer(synth.make_model(M=4, K=2, D=5, covariance="density"), "batch-diagonal-maximum-int")


@pytest.mark.parametrize("name", golden_names())
def test_golden_float_direct(eng, oracle, name):
    ms, z = load_golden(name)
    upload_float(eng, oracle_float_tables(ms))  # the reference build's tables (rsqrt-based isr)
    s, b = eng.score_host(z["x"])
    assert np.array_equal(bits(s), bits(z["max_score"]))  # reference operation order: bit-identical
    assert np.array_equal(b, z["max_best"])
    upload_float(eng, tables.float_tables(ms))  # the product's own IEEE tables
    s, b = eng.score_host(z["x"])
    assert_float_close(s, z["max_score"])
    upload_float(eng, oracle_float_tables(ms), mode=capi.MODE_SUM)
    s, b = eng.score_host(z["x"])
    assert_float_close(s, z["sum_score"])
    assert np.array_equal(b, z["sum_best"])


@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("mode", [capi.MODE_MAX, capi.MODE_SUM])
def test_golden_float_auto_kernel(eng, oracle, name, mode):
    ms, z = load_golden(name)
    upload_float(eng, tables.float_tables(ms), mode=mode, kernel=capi.KERNEL_AUTO)
    s, b = eng.score_host(z["x"])
    assert_float_close(s, z["max_score"] if mode == capi.MODE_MAX else z["sum_score"])
    t = oracle.float_tables(ms, exact_rsqrt=True)
    assert_best_consistent(ms, z["x"], b, z["max_best"] if mode == capi.MODE_MAX else z["sum_best"], oracle, t)


# ---------------------------------------------------------------------------------------------------
# seeded random models against the oracle
# ---------------------------------------------------------------------------------------------------
MODELS = [
    dict(M=257, K=16, D=45),
    dict(M=64, K=64, D=45, seed=2),
    dict(M=100, K=7, D=33, covariance="density", ragged=True, shared=True, zero_weight_frac=0.2, seed=3),
    dict(M=50, K=20, D=64, covariance="mixture", seed=4),
    dict(M=31, K=3, D=100, seed=5),
    dict(M=40, K=150, D=12, ragged=True, seed=6),
]


@pytest.mark.parametrize("kernel", ["auto", "tensor8", "dp4a"])
@pytest.mark.parametrize("T", [1, 130, 700])
def test_quantized_kernels_agree_with_oracle(eng, oracle, kernel, T, monkeypatch):
    """Pooled-covariance model: the integer tensor path (auto: 16 epilogue warps storing per lane quarter; tensor8:
    8 epilogue warps) and the forced dp4a kernel are all bit-exact."""
    if kernel == "dp4a":
        monkeypatch.setenv("GMM_QUANT_KERNEL", "dp4a")
    if kernel == "tensor8":
        monkeypatch.setenv("GMM_QTC_EPI_WARPS", "8")
    ms = synth.make_model(M=300, K=16, D=45, ragged=True, seed=11)
    x = synth.make_features(ms, T)
    x[-1] *= -40.0
    t, raw = oracle_quant_tables(ms)
    upload_quant(eng, t)
    assert eng.quant_kernel == (1 if kernel == "dp4a" else 2)
    r = eng.quant_score_host(x, want_qx=True)
    qx = oracle.simd_quantize_features(x, raw["isrs"])
    q, b = oracle.simd_score(ms.mix_offsets, raw, qx)
    assert np.array_equal(r["qx"], qx)
    assert np.array_equal(r["q"], q)
    assert np.array_equal(r["best"], b)
    assert np.array_equal(bits(r["score"]), bits(oracle.simd_final_score(q, raw["scaling_squared"])))


@pytest.mark.parametrize("warps", ["16", "8"])
@pytest.mark.parametrize("want", ["q+best", "score", "score+best", "q+best+score", "q", "generic"])
def test_quantized_output_sets_bit_exact(eng, oracle, want, warps, monkeypatch):
    """The integer tensor kernel is compiled per set of requested arrays (q + arg-min; score; score + arg-min; all;
    anything else and GMM_QTC_GENERIC take the run-time variant); the compiled sets produce the final f32 score by ONE
    f32 division instead of the reference's f64 expression (SimdFeatureScorer.cc:135-145): same bits required."""
    if want == "generic":
        monkeypatch.setenv("GMM_QTC_GENERIC", "1")
    if warps == "8":
        monkeypatch.setenv("GMM_QTC_EPI_WARPS", "8")
    ms = synth.make_model(M=301, K=16, D=45, seed=12)
    x = synth.make_features(ms, 391)
    x[7] *= 30.0
    t, raw = oracle_quant_tables(ms)
    upload_quant(eng, t)
    assert eng.quant_kernel == 2
    qx = oracle.simd_quantize_features(x, raw["isrs"])
    q, b = oracle.simd_score(ms.mix_offsets, raw, qx)
    sc = oracle.simd_final_score(q, raw["scaling_squared"])
    r = eng.quant_score_host(x, want_q="q" in want or want == "generic", want_best="best" in want or want == "generic",
                             want_score="score" in want or want == "generic")
    if r["q"] is not None:
        assert np.array_equal(r["q"], q)
    if r["best"] is not None:
        assert np.array_equal(r["best"], b)
    if r["score"] is not None:
        assert np.array_equal(bits(r["score"]), bits(sc))


@pytest.mark.parametrize("warps", ["16", "8"])
@pytest.mark.parametrize("K,ld", [(16, 304), (8, 308), (32, 304), (64, 312), (16, 301)])
def test_quantized_vector_stores_equal_scalar_stores(eng, oracle, K, ld, warps, monkeypatch):
    """Uniform tiles with 16-byte aligned row pieces take the vector store path of the integer tensor kernel (results
    in registers, 128-bit staging and stores); GMM_QTC_STORES=scalar forces the element-wise path.  Same bits in every
    requested array, nothing written beyond column M, and both equal the oracle."""
    if warps == "8":
        monkeypatch.setenv("GMM_QTC_EPI_WARPS", "8")
    ms = synth.make_model(M=301, K=K, D=45, seed=40 + K)
    T = 417
    x = synth.make_features(ms, T)
    x[3] *= 25.0
    t, raw = oracle_quant_tables(ms)
    upload_quant(eng, t)
    assert eng.quant_kernel == 2
    dx = torch.from_numpy(x).cuda()
    qx = oracle.simd_quantize_features(x, raw["isrs"])
    q, b = oracle.simd_score(ms.mix_offsets, raw, qx)
    sc = oracle.simd_final_score(q, raw["scaling_squared"])

    def run(want_q, want_b, want_s):
        dq = torch.full((T, ld), -5, dtype=torch.int32, device="cuda") if want_q else None
        db = torch.full((T, ld), -5, dtype=torch.int32, device="cuda") if want_b else None
        ds = torch.full((T, ld), -5.0, dtype=torch.float32, device="cuda") if want_s else None
        eng.quant_score_device(T, dx, dq, db, ds, ld=ld)
        eng.sync()
        return [None if a is None else a.cpu().numpy() for a in (dq, db, ds)]

    for want in ((True, True, False), (False, False, True), (False, True, True), (True, True, True)):
        monkeypatch.delenv("GMM_QTC_STORES", raising=False)
        vec = run(*want)
        monkeypatch.setenv("GMM_QTC_STORES", "scalar")
        sca = run(*want)
        for a, c, ref in zip(vec, sca, (q, b.astype(np.int32), sc)):
            if a is None:
                continue
            assert np.array_equal(bits(a), bits(c))
            assert np.array_equal(bits(a[:, :ms.M]), bits(ref))
            assert np.all(a[:, ms.M:] == -5)


@pytest.mark.parametrize("kw", MODELS)
@pytest.mark.parametrize("T", [1, 33, 300])
def test_quantized_bit_exact_vs_oracle(eng, oracle, kw, T):
    ms = synth.make_model(**kw)
    x = synth.make_features(ms, T)
    x[0] *= 50.0  # saturate some components
    t, raw = oracle_quant_tables(ms)
    upload_quant(eng, t)
    r = eng.quant_score_host(x, want_qx=True)
    qx = oracle.simd_quantize_features(x, raw["isrs"])
    q, b = oracle.simd_score(ms.mix_offsets, raw, qx)
    assert np.array_equal(r["qx"], qx)
    assert np.array_equal(r["q"], q)
    assert np.array_equal(r["best"], b)
    assert np.array_equal(bits(r["score"]), bits(oracle.simd_final_score(q, raw["scaling_squared"])))
    # the product's own (IEEE) table builder gives the tables of a non-fast-math reference build
    t2, raw2 = oracle_quant_tables(ms, exact_rsqrt=True)
    mine = tables.quant_tables(ms)
    upload_quant(eng, mine)
    r2 = eng.quant_score_host(x)
    qx2 = oracle.simd_quantize_features(x, raw2["isrs"])
    q2, b2 = oracle.simd_score(ms.mix_offsets, raw2, qx2)
    assert np.array_equal(r2["q"], q2) and np.array_equal(r2["best"], b2)


@pytest.mark.parametrize("kw", MODELS)
@pytest.mark.parametrize("T", [1, 33, 300])
def test_float_direct_bit_exact_vs_oracle(eng, oracle, kw, T):
    ms = synth.make_model(**kw)
    x = synth.make_features(ms, T)
    tab = oracle_float_tables(ms, mixture_weight_scale=0.9, gaussian_scale=1.2)
    upload_float(eng, tab)
    s, b = eng.score_host(x)
    so, bo = oracle.score_max(ms, x, tables=(tab["isr"], tab["log_norm"], tab["w2"]))
    assert np.array_equal(bits(s), bits(so))
    assert np.array_equal(b, bo)
    upload_float(eng, tab, mode=capi.MODE_SUM)
    s, b = eng.score_host(x)
    so, bo = oracle.score_sum(ms, x, tables=(tab["isr"], tab["log_norm"], tab["w2"]))
    assert_float_close(s, so)
    assert np.array_equal(b, bo)


@pytest.mark.parametrize("kw", MODELS)
@pytest.mark.parametrize("mode", [capi.MODE_MAX, capi.MODE_SUM])
def test_float_auto_kernel_vs_oracle(eng, oracle, kw, mode):
    ms = synth.make_model(**kw)
    x = synth.make_features(ms, 300)
    tab = tables.float_tables(ms)
    upload_float(eng, tab, mode=mode, kernel=capi.KERNEL_AUTO)
    s, b = eng.score_host(x)
    f = oracle.score_max if mode == capi.MODE_MAX else oracle.score_sum
    so, bo = f(ms, x, tables=(tab["isr"], tab["log_norm"], tab["w2"]))
    assert_float_close(s, so)
    assert_best_consistent(ms, x, b, bo, oracle, (tab["isr"], tab["log_norm"], tab["w2"]))


def test_posteriors(eng, oracle):
    import torch
    ms = synth.make_model(M=40, K=9, D=45, ragged=True, seed=8)
    x = synth.make_features(ms, 70)
    tab = oracle_float_tables(ms)
    upload_float(eng, tab, mode=capi.MODE_SUM)
    dx = torch.from_numpy(x).cuda()
    post = torch.empty((70, ms.G), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    eng.posteriors_device(70, dx, post)
    eng.sync()
    _, _, po = oracle.score_sum(ms, x, tables=(tab["isr"], tab["log_norm"], tab["w2"]), posterior=True)
    np.testing.assert_allclose(post.cpu().numpy(), po, rtol=1e-4, atol=1e-7)  # f32 exp / log of scores around D


def test_mixture_posteriors_host(eng, oracle):
    """gmm_mixture_posteriors_host (getDensityPosteriorProbabilities of the sum scorer for the mixture a frame is aligned
    to) against the oracle's posteriors: 1e-4 relative (f32 exp / log of scores around D)."""
    ms = synth.make_model(M=40, K=9, D=45, ragged=True, seed=8)
    T = 200
    x = synth.make_features(ms, T)
    tab = oracle_float_tables(ms)
    upload_float(eng, tab, mode=capi.MODE_SUM)
    rng = np.random.default_rng(4)
    mixture = rng.integers(0, ms.M, T).astype(np.uint32)
    mixture[5] = 0xFFFFFFFF
    max_k = int(np.max(np.diff(ms.mix_offsets)))
    post, score = eng.mixture_posteriors_host(x, mixture, max_k)
    so, _, po = oracle.score_sum(ms, x, tables=(tab["isr"], tab["log_norm"], tab["w2"]), posterior=True)
    for t in range(T):
        if mixture[t] == 0xFFFFFFFF:
            assert not post[t].any()
            continue
        a, b = int(ms.mix_offsets[mixture[t]]), int(ms.mix_offsets[mixture[t] + 1])
        np.testing.assert_allclose(post[t, :b - a], po[t, a:b], rtol=1e-4, atol=1e-7)
        assert not post[t, b - a:].any()
        assert abs(score[t] - so[t, mixture[t]]) <= 1e-5 * max(1.0, abs(so[t, mixture[t]]))
    upload_float(eng, tab, mode=capi.MODE_MAX)
    with pytest.raises(capi.GmmError):
        eng.mixture_posteriors_host(x, mixture, max_k)


# ---------------------------------------------------------------------------------------------------
# edge cases
# ---------------------------------------------------------------------------------------------------
def test_empty_batch_and_empty_mixture(eng, oracle):
    ms = synth.make_model(M=6, K=4, D=10, seed=9)
    # mixture 2 gets no densities at all
    offs = ms.mix_offsets.copy()
    offs[3:] -= (offs[3] - offs[2])
    keep = np.r_[0:ms.mix_offsets[2], ms.mix_offsets[3]:ms.G]
    from rasr_b200.gms import MixtureSetData
    ms2 = MixtureSetData(ms.D, offs, ms.mix_density[keep], ms.mix_logw[keep], ms.dens_mean, ms.dens_cov, ms.means,
                         ms.variances)
    x = synth.make_features(ms, 5)
    tab = oracle_float_tables(ms2)
    upload_float(eng, tab, kernel=capi.KERNEL_AUTO)
    s, b = eng.score_host(x)
    so, bo = oracle.score_max(ms2, x)
    assert np.array_equal(bits(s[:, 2]), bits(so[:, 2])) and (b[:, 2] == 0xFFFFFFFF).all()
    assert_float_close(s, so)
    s0, b0 = eng.score_host(np.zeros((0, ms.D), np.float32))
    assert s0.shape == (0, ms.M)
    t, raw = oracle_quant_tables(ms2)
    upload_quant(eng, t)
    r = eng.quant_score_host(x)
    q, bq = oracle.simd_score(ms2.mix_offsets, raw, oracle.simd_quantize_features(x, raw["isrs"]))
    assert np.array_equal(r["q"], q) and np.array_equal(r["best"], bq)


def test_argument_errors(eng):
    ms = synth.make_model(M=4, K=2, D=6)
    tab = tables.float_tables(ms)
    upload_float(eng, tab)
    with pytest.raises(ValueError):
        eng.score_host(np.zeros((3, 5), np.float32))
    with pytest.raises(capi.GmmError):
        eng.score_host(np.zeros((3, 6), np.float32), ld=2)
    bad = dict(tab)
    bad["cov_idx"] = tab["cov_idx"] + 5
    with pytest.raises(capi.GmmError):
        upload_float(eng, bad)
    e2 = capi.Engine(0)
    with pytest.raises(capi.GmmError):
        e2.D = 6
        e2.M = 4
        e2.score_host(np.zeros((3, 6), np.float32))
    e2.close()


def test_padded_leading_dimension(eng, oracle):
    ms = synth.make_model(M=37, K=8, D=20, seed=10)
    x = synth.make_features(ms, 19)
    upload_float(eng, oracle_float_tables(ms), kernel=capi.KERNEL_AUTO)
    s, b = eng.score_host(x, ld=64)
    so, bo = oracle.score_max(ms, x)
    assert_float_close(s[:, :37], so)


@pytest.mark.parametrize("K,ld", [(16, 308), (64, 302), (32, 304), (8, 320), (16, 301), (16, 306)])
def test_direct_row_piece_stores_equal_staged_stores(eng, oracle, K, ld, monkeypatch):
    """Maximum approximation without arg-min on uniform tiles: with 8- / 16-byte aligned row pieces (ld a multiple of
    2 / 4) the tensor kernel's lanes store their row pieces straight from registers; otherwise results are staged in
    shared memory and stored coalesced (GMM_TC_STORES=staged forces that).  Same bits, nothing written beyond column M."""
    ms = synth.make_model(M=301, K=K, D=45, seed=20 + K)
    T = 333
    x = synth.make_features(ms, T)
    upload_float(eng, tables.float_tables(ms), kernel=capi.KERNEL_TENSOR)
    dx = torch.from_numpy(x).cuda()

    def run():
        d = torch.full((T, ld), -123.0, dtype=torch.float32, device="cuda")
        eng.score_device(T, dx, d, None, ld=ld)
        eng.sync()
        return d.cpu().numpy()

    direct = run()  # 128-bit stores through the staging block where a lane holds >= 4 results, else lane stores
    monkeypatch.setenv("GMM_TC_STORES", "lane")
    lane = run()    # every lane writes its own row piece
    monkeypatch.setenv("GMM_TC_STORES", "staged")
    staged = run()
    assert np.array_equal(bits(direct), bits(staged)) and np.array_equal(bits(lane), bits(staged))
    assert np.all(direct[:, ms.M:] == -123.0)
    so, _ = oracle.score_max(ms, x)
    assert_float_close(direct[:, :ms.M], so)


@pytest.mark.parametrize("K,ld", [(16, 308), (64, 304), (32, 304), (8, 320), (16, 301)])
def test_full_sum_epilogue_variants_agree(eng, oracle, K, ld, monkeypatch):
    """Full-sum scorer on uniform tiles (sixteen epilogue warps, twenty-warp launch with setmaxnreg): parts without dead
    columns take an unpredicated pass and leave as row pieces from registers.  GMM_TC_SUM_FULL=0 forces the predicated
    pass with its skip vote, GMM_TC_STORES=staged the staged stores: the stores change no bit, the passes agree to
    rounding (different summation paths are NOT taken: same order, same terms -- identical bits expected as well), and
    all of them meet the oracle; nothing is written beyond column M."""
    ms = synth.make_model(M=301, K=K, D=45, seed=40 + K)
    T = 333
    x = synth.make_features(ms, T)
    upload_float(eng, tables.float_tables(ms), mode=capi.MODE_SUM, kernel=capi.KERNEL_TENSOR)
    dx = torch.from_numpy(x).cuda()

    def run():
        d = torch.full((T, ld), -123.0, dtype=torch.float32, device="cuda")
        eng.score_device(T, dx, d, None, ld=ld)
        eng.sync()
        return d.cpu().numpy()

    full = run()
    monkeypatch.setenv("GMM_TC_STORES", "staged")
    staged = run()
    monkeypatch.setenv("GMM_TC_SUM_FULL", "0")
    predicated = run()
    assert np.array_equal(bits(full), bits(staged))
    assert np.all(full[:, ms.M:] == -123.0) and np.all(predicated[:, ms.M:] == -123.0)
    so, _ = oracle.score_sum(ms, x)
    assert_float_close(full[:, :ms.M], so)
    assert_float_close(predicated[:, :ms.M], so)
    assert_float_close(full[:, :ms.M], predicated[:, :ms.M], rtol=2e-6)


def test_sixteen_warp_maximum_kernel_matches_eight(eng, oracle, monkeypatch):
    """GMM_TC_EPI_WARPS=16 selects the sixteen-warp epilogue for the maximum approximation too (twenty-warp launch,
    setmaxnreg): same bits as the default eight-warp kernel."""
    ms = synth.make_model(M=301, K=16, D=45, seed=61)
    x = synth.make_features(ms, 333)
    t = tables.float_tables(ms)
    upload_float(eng, t, kernel=capi.KERNEL_TENSOR)
    eight, _ = eng.score_host(x, want_best=False)
    monkeypatch.setenv("GMM_TC_EPI_WARPS", "16")
    upload_float(eng, t, kernel=capi.KERNEL_TENSOR)       # the epilogue width is chosen when the model is packed
    sixteen, _ = eng.score_host(x, want_best=False)
    assert np.array_equal(bits(eight), bits(sixteen))
    so, _ = oracle.score_max(ms, x)
    assert_float_close(sixteen, so)


# ---------------------------------------------------------------------------------------------------
# full-size configurations of BASELINE.json
# ---------------------------------------------------------------------------------------------------
def viterbi_path(scores, states):
    """Best path of a left-to-right HMM (loop / forward transitions, no skip; the Search::Aligner recursion of
    /root/reference/src/Search/Aligner.cc:206-239 with zero transition scores) through emission `states`."""
    T, S = scores.shape[0], len(states)
    em = scores[:, states].astype(np.float64)
    cost = np.full((T, S), np.inf)
    back = np.zeros((T, S), np.int32)
    cost[0, 0] = em[0, 0]
    for t in range(1, T):
        stay = cost[t - 1]
        move = np.r_[np.inf, cost[t - 1, :-1]]
        back[t] = (move < stay).astype(np.int32)
        cost[t] = np.minimum(stay, move) + em[t]
    path = [S - 1]
    for t in range(T - 1, 0, -1):
        path.append(path[-1] - back[t, path[-1]])
    return np.array(path[::-1])


def test_cfg1_full_size_float_and_alignment(eng, oracle):
    ms, x = synth.make_config("cfg1")  # 4501 x 16 x 45, T = 1000
    tab = oracle_float_tables(ms)
    so, bo = oracle.score_max(ms, x)
    upload_float(eng, tab, kernel=capi.KERNEL_DIRECT)
    s, b = eng.score_host(x)
    assert np.array_equal(bits(s), bits(so)) and np.array_equal(b, bo)
    upload_float(eng, tables.float_tables(ms), kernel=capi.KERNEL_AUTO)
    s2, b2 = eng.score_host(x)
    assert_float_close(s2, so)
    rng = np.random.default_rng(5)
    for seg in range(4):  # identical best-path alignments on the test segments
        states = rng.integers(0, ms.M, size=60)
        fr = slice(seg * 250, (seg + 1) * 250)
        assert np.array_equal(viterbi_path(s2[fr], states), viterbi_path(so[fr], states))


def test_cfg2_full_size_quantized(eng, oracle):
    ms, x = synth.make_config("cfg2")
    t, raw = oracle_quant_tables(ms)
    upload_quant(eng, t)
    r = eng.quant_score_host(x)
    q, b = oracle.simd_score(ms.mix_offsets, raw, oracle.simd_quantize_features(x, raw["isrs"]))
    assert np.array_equal(r["q"], q) and np.array_equal(r["best"], b)
    assert np.array_equal(bits(r["score"]), bits(oracle.simd_final_score(q, raw["scaling_squared"])))


def test_batches_beyond_one_launch(eng, oracle):
    """Long batches go out as several L2-sized launches (32768 frames float, 65536 quantized): rows on either side of
    every launch boundary must be the oracle's, and a frame's result must not depend on where the batch was cut."""
    ms = synth.make_model(M=260, K=16, D=45, seed=31)
    T = 70001
    x = synth.make_features(ms, T, seed=8)
    tab = tables.float_tables(ms)
    upload_float(eng, tab, kernel=capi.KERNEL_TENSOR)
    s, b = eng.score_host(x)
    rows = np.r_[0:4, 32760:32776, 65530:65544, T - 5:T]
    so, bo = oracle.score_max(ms, x[rows], tables=(tab["isr"], tab["log_norm"], tab["w2"]))
    assert_float_close(s[rows], so)
    s2, b2 = eng.score_host(x[40000:])               # other cut points: 40000 + 32768, ...
    assert np.array_equal(bits(s2), bits(s[40000:])) and np.array_equal(b2, b[40000:])
    t, raw = oracle_quant_tables(ms)
    upload_quant(eng, t)
    r = eng.quant_score_host(x)
    rows = np.r_[0:4, 65530:65544, T - 5:T]
    q, bq = oracle.simd_score(ms.mix_offsets, raw, oracle.simd_quantize_features(x[rows], raw["isrs"]))
    assert np.array_equal(r["q"][rows], q) and np.array_equal(r["best"][rows], bq)
    r2 = eng.quant_score_host(x[30000:])
    assert np.array_equal(r2["q"], r["q"][30000:]) and np.array_equal(r2["best"], r["best"][30000:])


@pytest.mark.parametrize("mode", [capi.MODE_MAX, capi.MODE_SUM])
def test_cfg3_large_model_properties(eng, oracle, mode):
    """12001 x 64: oracle on a frame sample; batch-split invariance and frame-permutation equivariance at T=10k."""
    ms, x = synth.make_config("cfg3", T=10000)
    tab = tables.float_tables(ms)
    upload_float(eng, tab, mode=mode, kernel=capi.KERNEL_AUTO)
    s, _ = eng.score_host(x, want_best=False)
    idx = np.random.default_rng(1).choice(10000, 48, replace=False)
    f = oracle.score_max if mode == capi.MODE_MAX else oracle.score_sum
    so, _ = f(ms, x[idx], tables=(tab["isr"], tab["log_norm"], tab["w2"]))
    assert_float_close(s[idx], so)
    perm = np.random.default_rng(2).permutation(10000)[:3000]
    sp, _ = eng.score_host(x[perm], want_best=False)
    assert np.array_equal(bits(sp), bits(s[perm]))  # a frame's scores do not depend on its batch neighbours
    s1, _ = eng.score_host(x[:777], want_best=False)
    assert np.array_equal(bits(s1), bits(s[:777]))


def test_feature_scorer_interface_on_gpu(oracle):
    ms = synth.make_model(M=120, K=8, D=45, seed=12)
    x = synth.make_features(ms, 50)
    so, bo = oracle.score_max(ms, x)
    fs = feature_scorer.GpuFeatureScorer(ms, "diagonal-maximum", buffer_size=8, float_tables=oracle_float_tables(ms),
                                         kernel=capi.KERNEL_DIRECT)
    got = []
    for f in x:
        if not fs.bufferFilled():
            fs.addFeature(f)
        else:
            got.append(fs.getScorer(f))
            got[-1].score(0)
    while not fs.bufferEmpty():
        got.append(fs.flush())
    assert len(got) == 50
    for t in (0, 7, 8, 49):
        assert got[t].score(5) == so[t, 5] and got[t].bestDensity(5) == bo[t, 5]
    t, raw = oracle_quant_tables(ms)
    fq = feature_scorer.GpuFeatureScorer(ms, "SIMD-diagonal-maximum", quant_tables=t)
    for f in x:
        fq.addFeature(f)
    q, bq = oracle.simd_score(ms.mix_offsets, raw, oracle.simd_quantize_features(x, raw["isrs"]))
    sc = oracle.simd_final_score(q, raw["scaling_squared"])
    for t_ in range(50):
        c = fq.flush()
        assert c.score(3) == sc[t_, 3] and c.bestDensity(3) == bq[t_, 3]


# ---------------------------------------------------------------------------------------------------
# host entry points: chunked pipeline and the asynchronous pair
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("ld", [257, 260])
def test_chunked_host_call_equals_one_chunk(eng, oracle, ld, monkeypatch):
    """gmm_score_host scores long batches in chunks and copies finished chunks back on a second stream; the result is
    the single-chunk call's, bit for bit (rows and arg-min), for an unaligned and an aligned row pitch."""
    ms = synth.make_model(M=257, K=16, D=45, seed=31)
    T = 9000
    x = synth.make_features(ms, T)
    upload_float(eng, tables.float_tables(ms), kernel=capi.KERNEL_AUTO)
    monkeypatch.setenv("GMM_HOST_CHUNKS", "1")
    s1, b1 = eng.score_host(x, ld=ld)
    monkeypatch.setenv("GMM_HOST_CHUNKS", "5")
    s5, b5 = eng.score_host(x, ld=ld)
    monkeypatch.delenv("GMM_HOST_CHUNKS")
    sd, bd = eng.score_host(x, ld=ld)      # default plan (2 chunks at this size)
    for s, b in ((s5, b5), (sd, bd)):
        assert np.array_equal(bits(s[:, :ms.M]), bits(s1[:, :ms.M])) and np.array_equal(b[:, :ms.M], b1[:, :ms.M])
    so, _ = oracle.score_max(ms, x[:300])
    assert_float_close(s1[:300, :ms.M], so)


def test_async_pair_overlaps_calls_and_matches_blocking_call(eng):
    """gmm_score_async / gmm_wait: three calls enqueued back to back on page-locked buffers (the third
    reuses the first one's lane), waited for out of order; each equals the blocking call on the same frames."""
    ms = synth.make_model(M=301, K=16, D=45, seed=32)
    upload_float(eng, tables.float_tables(ms), kernel=capi.KERNEL_AUTO)
    T, ld = 1500, 304
    xs = [torch.from_numpy(synth.make_features(ms, T, seed=40 + i)).pin_memory() for i in range(3)]
    outs = [torch.full((T, ld), -1.0, dtype=torch.float32).pin_memory() for _ in range(3)]
    bests = [torch.zeros((T, ld), dtype=torch.int32).pin_memory() for _ in range(3)]
    toks = [eng.score_async(xs[i].numpy(), outs[i].numpy(), bests[i].numpy(), ld=ld) for i in range(3)]
    assert toks == sorted(toks) and len(set(toks)) == 3
    for i in (2, 0, 1):
        eng.wait(toks[i])
    eng.sync()
    for i in range(3):
        s, b = eng.score_host(xs[i].numpy(), ld=ld)
        assert np.array_equal(bits(outs[i].numpy()[:, :ms.M]), bits(s[:, :ms.M]))
        assert np.array_equal(bests[i].numpy().view(np.uint32)[:, :ms.M], b[:, :ms.M])
    with pytest.raises(capi.GmmError):
        eng.wait(10 ** 9)
