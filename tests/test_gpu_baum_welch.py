"""GPU parity of the Baum-Welch alignment consumer (gmm_baum_welch_device / gmm_score_baum_welch_host /
gmm_score_baum_welch_accumulate_host) against the oracle's f64 forward-backward restatement of Search::Aligner's
baum-welch mode.  Posteriors are probabilities in [0, 1]: tolerance 1e-4 absolute (the device keeps f32 potentials
relative to per-frame offsets summed in f64); total scores 1e-5 relative."""
import numpy as np
import pytest
import torch

from rasr_b200 import capi, synth, tables
from rasr_b200.align import AlignmentAutomata, linear_hmm

pytestmark = pytest.mark.gpu
FLT_MAX = float(np.finfo(np.float32).max)
TOL = 1e-4


@pytest.fixture(scope="module")
def eng():
    e = capi.Engine(0)
    yield e
    e.close()


def random_graph(rng, S, M):
    states = []
    for s in range(S):
        states.append([(int(rng.integers(0, S)), int(rng.integers(0, M)), float(rng.uniform(0, 5)))
                       for _ in range(int(rng.integers(0, 5)))])
    final = {int(s): float(rng.uniform(0, 3)) for s in rng.choice(S, size=max(1, S // 4), replace=False)}
    return states, int(rng.integers(0, S)), final


def check(eng, oracle, fo, scores, au):
    lo, lst = eng.emission_lists(au, scores.shape[1])
    ld_w = max(1, int(np.max(np.diff(lo)))) + 3
    d = torch.from_numpy(scores).cuda()
    w = torch.full((scores.shape[0], ld_w), -1.0, device="cuda")
    total = eng.baum_welch_device(fo, d, au, w)
    w = w.cpu().numpy()
    wo, lists, to = oracle.baum_welch_align(fo, scores, au)
    for i in range(len(fo) - 1):
        assert np.array_equal(lst[lo[i]:lo[i + 1]], lists[i])
        rows = slice(int(fo[i]), int(fo[i + 1]))
        n = len(lists[i])
        if to[i] == FLT_MAX:
            assert total[i] == FLT_MAX and not w[rows, :n].any()
            continue
        assert abs(total[i] - to[i]) <= 1e-5 * max(1.0, abs(to[i]))
        assert np.max(np.abs(w[rows, :n] - wo[i]), initial=0.0) <= TOL
        if rows.stop > rows.start and n:
            np.testing.assert_allclose(w[rows, :n].sum(axis=1), 1.0, atol=TOL)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_linear_automata_ragged_segments(eng, oracle, seed):
    rng = np.random.default_rng(seed)
    M = 500
    lens = [int(n) for n in rng.integers(1, 300, 9)] + [0, 1000]
    fo = np.cumsum([0] + lens).astype(np.uint32)
    scores = rng.uniform(20, 120, (int(fo[-1]), M)).astype(np.float32)
    segs = [linear_hmm(rng.integers(0, M, max(1, n // int(rng.integers(2, 6)))), loop=3.0, forward=0.0, skip=30.0)
            for n in lens]
    check(eng, oracle, fo, scores, AlignmentAutomata.from_arc_lists(segs))


@pytest.mark.parametrize("seed", [3, 4])
def test_general_graphs_and_unreachable_segments(eng, oracle, seed):
    rng = np.random.default_rng(seed)
    M = 60
    lens = [int(n) for n in rng.integers(1, 60, 14)]
    fo = np.cumsum([0] + lens).astype(np.uint32)
    scores = rng.uniform(0, 30, (int(fo[-1]), M)).astype(np.float32)
    segs = [random_graph(rng, int(rng.integers(1, 40)), M) for _ in lens]
    segs[3] = ([[(1, 0, 0.0)], []], 0, {1: 0.0})        # reaches its final state only with exactly one frame
    check(eng, oracle, fo, scores, AlignmentAutomata.from_arc_lists(segs))


def test_long_segment_keeps_precision(eng, oracle):
    """4000 frames: true potentials reach 3e5, where f32 has an ulp of 0.03 -- the relative potentials must not."""
    rng = np.random.default_rng(9)
    M, T = 300, 4000
    fo = np.array([0, T], np.uint32)
    scores = rng.uniform(40, 120, (T, M)).astype(np.float32)
    au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, M, 900), loop=3.0, forward=0.0, skip=30.0)])
    check(eng, oracle, fo, scores, au)


def test_features_to_posteriors_and_training_pass(eng, oracle):
    """Host features in: posterior rows over compact AND full score blocks agree with the oracle run on this engine's
    own scores; the training pass accumulates every frame with total weight 1 (the posteriors of a frame sum to 1)."""
    import os
    ms = synth.make_model(M=900, K=16, D=45, seed=21)
    rng = np.random.default_rng(2)
    lens = [250, 0, 130, 1, 400]
    fo = np.cumsum([0] + lens).astype(np.uint32)
    x = synth.make_features(ms, int(fo[-1]), seed=6)
    segs = [linear_hmm(rng.integers(0, ms.M, n), loop=3.0, forward=0.0, skip=30.0) for n in (60, 3, 40, 1, 120)]
    au = AlignmentAutomata.from_arc_lists(segs)
    t = tables.float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX,
                     kernel=capi.KERNEL_TENSOR)
    s, _ = eng.score_host(x, want_best=False)
    wo, lists, to = oracle.baum_welch_align(fo, s, au)
    got = {}
    for flag in ("0", "1"):
        os.environ["GMM_ALIGN_SUBSET"] = flag
        try:
            got[flag] = eng.score_baum_welch_host(fo, x, au)
        finally:
            del os.environ["GMM_ALIGN_SUBSET"]
    for flag in ("0", "1"):
        w, total = got[flag]
        for i in range(len(lens)):
            rows = slice(int(fo[i]), int(fo[i + 1]))
            n = len(lists[i])
            assert np.max(np.abs(w[rows, :n] - wo[i]), initial=0.0) <= TOL
            assert abs(total[i] - to[i]) <= 1e-5 * max(1.0, abs(to[i]))
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    _, total = eng.score_baum_welch_host(fo, x, au, accumulate=True)
    acc = eng.read_accumulators(ms.variances.shape[0])
    reached = sum(n for n, tt in zip(lens, to) if tt != FLT_MAX)
    np.testing.assert_allclose(np.sum(acc["mix_weight"]), reached, rtol=1e-4)
    np.testing.assert_allclose(total, to, rtol=1e-5)
    # per-mixture mass: sum over frames of the posterior of that emission
    want = np.zeros(ms.M)
    for i in range(len(lens)):
        for j, e in enumerate(lists[i]):
            want[int(e)] += float(np.sum(wo[i][:, j]))
    have = np.add.reduceat(acc["mix_weight"], ms.mix_offsets[:-1].astype(np.int64))
    np.testing.assert_allclose(have, want, atol=2e-3)
