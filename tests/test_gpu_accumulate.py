"""GPU parity of the training E-step (gmm_accumulate_*) against the oracle: the chosen
density of every aligned frame bit-exact, the unweighted counts exact, the f64 sums within 1e-12 relative (they are
added with atomics, i.e. in a different order than the reference's frame order)."""
import numpy as np
import pytest
import torch

from conftest import oracle_float_tables
from rasr_b200 import capi, synth, tables
from rasr_b200.align import AlignmentAutomata, linear_hmm

pytestmark = pytest.mark.gpu
SKIP = 0xFFFFFFFF
SUM_RTOL = 1e-12


@pytest.fixture(scope="module")
def eng():
    e = capi.Engine(0)
    yield e
    e.close()


def upload(eng, ms, mode, kernel=capi.KERNEL_DIRECT):
    t = oracle_float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=mode, kernel=kernel)
    return (t["isr"], t["log_norm"], t["w2"])


def assert_accumulators(got, want, exact_weights):
    for k in ("mean_sum", "cov_sum"):
        # relative to the size of the terms added (a mean sum can cancel to far below its terms)
        scale = np.abs(want[k]) + np.abs(want[k]).max()
        assert np.max(np.abs(got[k] - want[k]) / scale) <= SUM_RTOL, k
    for k in ("mean_weight", "cov_weight", "mix_weight"):
        if exact_weights:
            assert np.array_equal(got[k], want[k]), k
        else:
            np.testing.assert_allclose(got[k], want[k], rtol=SUM_RTOL, atol=0)


@pytest.mark.parametrize("covariance,D,K", [("pooled", 45, 16), ("density", 17, 5), ("mixture", 64, 40), ("pooled", 100, 3)])
@pytest.mark.parametrize("weighted", [False, True])
def test_viterbi_accumulation_matches_oracle(eng, oracle, covariance, D, K, weighted):
    ms = synth.make_model(M=97, K=K, D=D, seed=D, covariance=covariance)
    T = 3001
    x = synth.make_features(ms, T, seed=5)
    rng = np.random.default_rng(D + K)
    mixture = rng.integers(0, ms.M, T).astype(np.uint32)
    mixture[rng.random(T) < 0.05] = SKIP
    weight = rng.uniform(0.1, 2.0, T) if weighted else None
    tab = upload(eng, ms, capi.MODE_MAX)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    best = eng.accumulate_host(x, mixture, weight)
    want = oracle.accumulate(ms, x, mixture, weight, tables=tab)
    keep = mixture != SKIP
    assert np.array_equal(best[keep], want["best"][keep])
    got = eng.read_accumulators(ms.n_cov)
    assert_accumulators(got, want, exact_weights=not weighted)
    # a second batch adds on top; reset clears
    eng.accumulate_host(x, mixture, weight, want_best=False)
    got2 = eng.read_accumulators(ms.n_cov)
    np.testing.assert_allclose(got2["mix_weight"], 2 * want["mix_weight"], rtol=1e-12)
    np.testing.assert_allclose(got2["mean_sum"], 2 * want["mean_sum"], rtol=1e-11, atol=1e-9)
    eng.reset_accumulators()
    assert not eng.read_accumulators(ms.n_cov)["cov_sum"].any()


def test_sum_scorer_best_density_and_posterior_weights(eng, oracle):
    ms = synth.make_model(M=50, K=8, D=20, seed=12, covariance="density")
    T = 2000
    x = synth.make_features(ms, T, seed=6)
    rng = np.random.default_rng(3)
    mixture = rng.integers(0, ms.M, T).astype(np.uint32)
    weight = rng.uniform(0.5, 1.5, T)
    tab = upload(eng, ms, capi.MODE_SUM)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means, viterbi=True)
    best = eng.accumulate_host(x, mixture)
    want = oracle.accumulate(ms, x, mixture, sum_scorer=True, tables=tab)
    assert np.array_equal(best, want["best"])
    assert_accumulators(eng.read_accumulators(ms.n_cov), want, exact_weights=True)
    # posterior-weighted: device expf/logf differ from libm in the last bits -> 1e-5 relative on the weights
    upload(eng, ms, capi.MODE_SUM)
    thr = float(np.finfo(np.float32).eps)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means, viterbi=False, weight_threshold=thr)
    eng.accumulate_host(x, mixture, weight, want_best=False)
    want = oracle.accumulate(ms, x, mixture, weight, sum_scorer=True, viterbi=False, weight_threshold=thr, tables=tab)
    got = eng.read_accumulators(ms.n_cov)
    np.testing.assert_allclose(got["mix_weight"].sum(), weight.sum(), rtol=1e-5)
    for k in want:
        if k != "best":
            np.testing.assert_allclose(got[k], want[k], rtol=2e-5, atol=1e-4 * np.abs(want[k]).max())


def test_accumulator_errors(eng):
    ms = synth.make_model(M=5, K=2, D=8, seed=1)
    upload(eng, ms, capi.MODE_MAX)
    with pytest.raises(capi.GmmError, match="no float model or no accumulators"):
        eng.accumulate_host(np.zeros((1, 8), np.float32), np.zeros(1, np.uint32))
    with pytest.raises(capi.GmmError, match="sum scorer"):
        eng.create_accumulators(tables.mean_index(ms), ms.n_means, viterbi=False)
    with pytest.raises(capi.GmmError, match="out of range"):
        eng.create_accumulators(tables.mean_index(ms), 1)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    with pytest.raises(capi.GmmError, match="mixture index out of range"):
        eng.accumulate_host(np.zeros((1, 8), np.float32), np.array([5], np.uint32))
    with pytest.raises(capi.GmmError, match="already has accumulators"):
        eng.create_accumulators(tables.mean_index(ms), ms.n_means)


def test_align_then_accumulate_on_device(eng, oracle):
    """The training pass: features -> scores -> Viterbi path -> E-step, with the path read back only to check."""
    ms, x = synth.make_config("cfg1")
    rng = np.random.default_rng(4)
    fo = np.array([0, 300, 300, 640, 1000], np.uint32)
    segs = [linear_hmm(rng.integers(0, ms.M, n), loop=3.0, forward=0.0, skip=None) for n in (70, 2, 90, 500)]
    au = AlignmentAutomata.from_arc_lists(segs)      # the last automaton cannot finish in 360 frames: skipped
    tab = upload(eng, ms, capi.MODE_MAX, kernel=capi.KERNEL_AUTO)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    em, st, sc = eng.score_align_host(fo, x, au)
    assert np.all(em[640:] == SKIP) and np.all(em[:640] != SKIP)
    dx = torch.from_numpy(x).cuda()
    dm = torch.from_numpy(em.view(np.int32)).cuda()
    eng.accumulate_device(x.shape[0], dx, dm)
    got = eng.read_accumulators(ms.n_cov)
    want = oracle.accumulate(ms, x, em, tables=tab)
    assert got["mix_weight"].sum() == 640
    assert_accumulators(got, want, exact_weights=True)
    # the same as one call, with and without the paths coming back
    eng.reset_accumulators()
    em2, st2, sc2 = eng.score_align_host(fo, x, au, accumulate=True)
    assert np.array_equal(em2, em) and np.array_equal(st2, st) and np.array_equal(sc2, sc)
    assert_accumulators(eng.read_accumulators(ms.n_cov), want, exact_weights=True)
    eng.reset_accumulators()
    assert eng.score_align_host(fo, x, au, accumulate=True, want_paths=False) == (None, None, None)
    assert_accumulators(eng.read_accumulators(ms.n_cov), want, exact_weights=True)


def test_training_iteration_through_mixset_file(eng, oracle, tmp_path):
    """GPU E-step -> MIXSET accumulator file -> M-step -> text model -> upload: the model the reference's own
    tools would produce from the oracle's accumulators, to f32 rounding."""
    from rasr_b200 import model_io
    ms = synth.make_model(M=40, K=6, D=24, seed=21, covariance="pooled")
    T = 20000
    x = synth.make_features(ms, T, seed=8)
    mixture = np.random.default_rng(8).integers(0, ms.M, T).astype(np.uint32)
    tab = upload(eng, ms, capi.MODE_MAX)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    eng.accumulate_host(x, mixture, want_best=False)
    est = model_io.MixtureSetEstimatorData.from_accumulators(ms, eng.read_accumulators(ms.n_cov))
    model_io.write_mixset(est, tmp_path / "acc.mix")
    new = model_io.read_mixset(tmp_path / "acc.mix").estimate()
    ref = model_io.MixtureSetEstimatorData.from_accumulators(ms, oracle.accumulate(ms, x, mixture, tables=tab)).estimate()
    assert np.array_equal(new.mix_offsets, ref.mix_offsets) and np.array_equal(new.mix_density, ref.mix_density)
    assert np.array_equal(new.mix_logw, ref.mix_logw)                 # counts are exact
    np.testing.assert_allclose(new.means, ref.means, rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(new.variances, ref.variances, rtol=1e-6)
    model_io.write_mixture_set_text_exact(new, tmp_path / "new.pms.gz")
    loaded = model_io.read_mixture_set_text(tmp_path / "new.pms.gz")
    t = oracle_float_tables(loaded)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX,
                     kernel=capi.KERNEL_DIRECT)
    s, b = eng.score_host(x[:64])
    so, bo = oracle.score_max(loaded, x[:64])
    assert np.array_equal(s.view(np.uint32), so.view(np.uint32)) and np.array_equal(b, bo)


def test_export_import_block_layout(eng, oracle):
    """The packed device block (what ranks all-reduce) is the concatenation read_accumulators returns."""
    from rasr_b200 import sharded
    ms = synth.make_model(M=20, K=3, D=10, seed=2, covariance="mixture")
    x = synth.make_features(ms, 500, seed=1)
    mixture = np.random.default_rng(1).integers(0, ms.M, 500).astype(np.uint32)
    upload(eng, ms, capi.MODE_MAX)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    eng.accumulate_host(x, mixture, want_best=False)
    acc = eng.read_accumulators(ms.n_cov)
    n = eng.accumulator_size
    assert n == ms.n_means * ms.D + ms.n_means + ms.n_cov * ms.D + ms.n_cov + ms.G
    block = torch.empty(n, dtype=torch.float64, device="cuda")
    eng.export_accumulators(block)
    eng.sync()
    assert np.array_equal(block.cpu().numpy(), sharded.pack_accumulators(acc))
    eng.import_accumulators(block * 2)
    eng.sync()
    twice = eng.read_accumulators(ms.n_cov)
    assert all(np.array_equal(twice[k], 2 * acc[k]) for k in acc)
    back = sharded.unpack_accumulators(block.cpu().numpy(), ms.n_means, ms.n_cov, ms.D, ms.G)
    assert all(np.array_equal(back[k], acc[k]) for k in acc)


# ---------------------------------------------------------------------------------------------------
# the reference's own accumulator files (tests/golden/est_*.npz, written by oracle/_ref/ref_estimator)
# ---------------------------------------------------------------------------------------------------
from conftest import ESTIMATOR_CASES, estimator_golden_names, load_golden  # noqa: E402
from rasr_b200 import model_io  # noqa: E402


@pytest.mark.parametrize("case", sorted(ESTIMATOR_CASES))
@pytest.mark.parametrize("name", estimator_golden_names())
def test_e_step_against_reference_accumulator_files(tmp_path, eng, name, case):
    """gmm_accumulate_host -> MIXSET file -> M-step against AbstractMixtureSetEstimator::accumulate / write / estimate
    of the unmodified reference (AbstractMixtureSetEstimator.cc:117-147, 482-506, 299-335)."""
    ms, z = load_golden(name)
    scorer, viterbi, weighted = ESTIMATOR_CASES[case]
    upload(eng, ms, capi.MODE_SUM if scorer == "sum" else capi.MODE_MAX)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means, viterbi=viterbi)
    best = eng.accumulate_host(z["x"], z["mixture"], z["weight"] if weighted else None)
    (tmp_path / "ref.mixset").write_bytes(bytes(z["mixset_" + case]))
    ref = model_io.read_mixset(tmp_path / "ref.mixset")
    got = model_io.MixtureSetEstimatorData.from_accumulators(ms, eng.read_accumulators(ms.n_cov))
    model_io.write_mixset(got, tmp_path / "gpu.mixset")          # renumbers the estimators as the reference's writer does
    got = model_io.read_mixset(tmp_path / "gpu.mixset")
    assert got.equal_topology(ref)
    if viterbi:
        live = z["mixture"] != SKIP
        assert np.array_equal(best[live], z["best_" + case][live])
    # viterbi: same terms, added by atomics in another order (1e-12); posterior weights: device expf / logf (2e-5)
    rtol = SUM_RTOL if viterbi else 2e-5
    for f in ("mix_weight", "mean_weight", "cov_weight", "mean_sum", "cov_sum"):
        a, b = getattr(got, f), getattr(ref, f)
        assert np.max(np.abs(a - b)) <= rtol * np.max(np.abs(b)), f
    if viterbi and not weighted:
        assert np.array_equal(got.mix_weight, ref.mix_weight)    # counts are exact
        # the M-step of the GPU's accumulators, printed like the reference prints its own: identical text
        model_io.write_mixture_set_text(got.estimate(), tmp_path / "est.pms")
        assert (tmp_path / "est.pms").read_bytes() == bytes(z["est_" + case])
