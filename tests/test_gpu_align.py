"""GPU parity of the device-resident alignment consumer (gmm_align_device / gmm_score_align_host, SURVEY.md 8(f)
rank 1) against the oracle's restatement of Search::Aligner's Viterbi search space: bit-exact paths and scores,
including exact f32 ties, which must be resolved in the reference's hypothesis order."""
import numpy as np
import pytest
import torch

from conftest import oracle_float_tables
from rasr_b200 import capi, synth, tables
from rasr_b200.align import AlignmentAutomata, linear_hmm

pytestmark = pytest.mark.gpu
FLT_MAX = np.finfo(np.float32).max


@pytest.fixture(scope="module")
def eng():
    e = capi.Engine(0)
    yield e
    e.close()


@pytest.fixture(params=["ranked", "generic"])
def kernel(request, monkeypatch):
    """Both device kernels: the rank-carrying one (default) and the order-free one with ordered replay on ties."""
    if request.param == "generic":
        monkeypatch.setenv("GMM_ALIGN_KERNEL", "generic")
    else:
        monkeypatch.delenv("GMM_ALIGN_KERNEL", raising=False)
    return request.param


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def random_graph(rng, S, M, integer_weights):
    """General automaton: random out-degree, random targets, arcs in random order, several final states."""
    states = []
    for s in range(S):
        arcs = []
        for _ in range(int(rng.integers(0, 5))):
            w = float(rng.integers(0, 3)) if integer_weights else float(rng.uniform(0, 5))
            arcs.append((int(rng.integers(0, S)), int(rng.integers(0, M)), w))
        states.append(arcs)
    final = {int(s): (float(rng.integers(0, 2)) if integer_weights else float(rng.uniform(0, 3)))
             for s in rng.choice(S, size=max(1, S // 4), replace=False)}
    return states, int(rng.integers(0, S)), final


def check(eng, oracle, fo, scores, au):
    d = torch.from_numpy(scores).cuda()
    n0 = eng.align_ordered_count
    em, st, sc = eng.align_device(fo, d, au, ld=scores.shape[1])
    emo, sto, sco = oracle.viterbi_align(fo, scores, au)
    assert np.array_equal(bits(sc), bits(sco))
    assert np.array_equal(em, emo) and np.array_equal(st, sto)
    em2, st2, sc2 = eng.align_device(fo, d, au, ld=scores.shape[1], want_states=False)
    assert st2 is None and np.array_equal(em2, emo)
    return eng.align_ordered_count - n0


@pytest.mark.parametrize("seed", range(4))
def test_linear_automata_ragged_segments(eng, oracle, kernel, seed):
    rng = np.random.default_rng(seed)
    M, segs, fo = 300, [], [0]
    for i in range(int(rng.integers(40, 400))):
        n = int(rng.integers(1, 120))
        T = 0 if i % 37 == 5 else int(rng.integers(max(1, n // 3), 3 * n + 5))   # some unreachable, some empty
        segs.append(linear_hmm(rng.integers(0, M, n), loop=rng.uniform(0, 4), forward=rng.uniform(0, 2),
                               skip=rng.uniform(2, 8) if i % 3 else None))
        fo.append(fo[-1] + T)
    scores = rng.uniform(0, 80, size=(fo[-1], M)).astype(np.float32)
    ordered = check(eng, oracle, np.array(fo, np.uint32), scores, AlignmentAutomata.from_arc_lists(segs))
    # f32 path scores in the thousands tie often: the generic kernel replays many segments, the ranked one none
    assert ordered == 0 or kernel == "generic"


@pytest.mark.parametrize("seed", range(4))
def test_general_graphs_with_ties(eng, oracle, kernel, seed):
    """Integer scores and weights: ties everywhere, states not numbered in activation order."""
    rng = np.random.default_rng(100 + seed)
    M, segs, fo = 17, [], [0]
    for _ in range(60):
        segs.append(random_graph(rng, int(rng.integers(1, 70)), M, integer_weights=True))
        fo.append(fo[-1] + int(rng.integers(0, 40)))
    scores = rng.integers(0, 3, size=(fo[-1], M)).astype(np.float32)
    ordered = check(eng, oracle, np.array(fo, np.uint32), scores, AlignmentAutomata.from_arc_lists(segs))
    assert (ordered >= 20) if kernel == "generic" else (ordered == 0)


def test_general_graphs_continuous_scores(eng, oracle, kernel):
    rng = np.random.default_rng(7)
    M, segs, fo = 64, [], [0]
    for _ in range(200):
        segs.append(random_graph(rng, int(rng.integers(1, 400)), M, integer_weights=False))
        fo.append(fo[-1] + int(rng.integers(0, 60)))
    scores = rng.uniform(0, 30, size=(fo[-1], M)).astype(np.float32)
    scores[rng.random(scores.shape) < 0.02] = np.inf      # unusable emissions
    check(eng, oracle, np.array(fo, np.uint32), scores, AlignmentAutomata.from_arc_lists(segs))


def test_tie_rules_of_the_reference(eng, oracle, kernel):
    states = [[(1, 3, 0.0), (2, 4, 0.0)], [(3, 5, 0.0)], [(3, 6, 0.0)], []]
    a = AlignmentAutomata.from_arc_lists([(states, 0, {3: 0.0}), ([[(2, 1, 0.0), (1, 0, 0.0)], [], []], 0,
                                                                  {1: 0.0, 2: 0.0})])
    scores = np.ones((3, 8), np.float32)
    em, st, sc = eng.align_device([0, 2, 3], torch.from_numpy(scores).cuda(), a, ld=8)
    assert list(em) == [4, 6, 1] and list(st) == [2, 3, 2] and list(sc) == [2.0, 1.0]
    assert check(eng, oracle, np.array([0, 2, 3], np.uint32), scores, a) == (4 if kernel == "generic" else 0)


def test_large_automaton_and_long_segment(eng, oracle, kernel):
    rng = np.random.default_rng(3)
    M = 4501
    segs = [linear_hmm(rng.integers(0, M, 3000), loop=1.0, forward=0.5, skip=4.0),
            linear_hmm(rng.integers(0, M, 7), loop=0.1, forward=0.5, skip=None)]
    fo = np.array([0, 4000, 4600], np.uint32)
    scores = rng.uniform(0, 100, size=(4600, M)).astype(np.float32)
    check(eng, oracle, fo, scores, AlignmentAutomata.from_arc_lists(segs))


def test_wide_fan_out_and_oscillating_lists(eng, oracle, kernel):
    """An entry state with hundreds of arcs (wide arrival keys) and a bipartite automaton whose hypothesis list
    never becomes a fixed point (the ranking is recomputed every frame)."""
    rng = np.random.default_rng(21)
    M = 50
    fan = [[(int(t), int(rng.integers(0, M)), float(rng.integers(0, 2))) for t in rng.permutation(300) + 1]]
    fan += [[(int(rng.integers(1, 301)), int(rng.integers(0, M)), 0.0), (s, int(rng.integers(0, M)), 1.0)]
            for s in range(1, 301)]
    bip = [[(1, 3, 0.0), (3, 4, 0.0)], [(0, 5, 0.0), (2, 6, 1.0)], [(1, 7, 0.0), (3, 8, 0.0)], [(2, 9, 0.0), (0, 1, 0.0)]]
    segs = [(fan, 0, {7: 0.0, 9: 0.0, 200: 1.0}), (bip, 0, {0: 0.0, 1: 0.0, 2: 0.0, 3: 0.0})]
    fo = np.array([0, 40, 75], np.uint32)
    scores = rng.integers(0, 4, size=(75, M)).astype(np.float32)
    check(eng, oracle, fo, scores, AlignmentAutomata.from_arc_lists(segs))


def test_argument_errors(eng):
    a = AlignmentAutomata.from_arc_lists([linear_hmm([1, 2])])
    d = torch.zeros((2, 4), device="cuda")
    bad = AlignmentAutomata.from_arc_lists([linear_hmm([1, 9])])
    with pytest.raises(capi.GmmError, match="emission index"):
        eng.align_device([0, 2], d, bad, ld=4)
    a.arc_target[0] = 77
    with pytest.raises(capi.GmmError, match="arc target"):
        eng.align_device([0, 2], d, a, ld=4)
    with pytest.raises(ValueError):
        eng.align_device([0, 1, 2], d, a, ld=4)
    em, st, sc = eng.align_device(np.zeros(1, np.uint32), d, AlignmentAutomata.from_arc_lists([]), ld=4)
    assert em.size == 0 and sc.size == 0


def test_score_and_align_from_host_features(eng, oracle):
    """Features in, paths out: the path must be the oracle's Viterbi path over the scores this engine computes
    (bit-exact), and over the oracle's own scores (north_star: identical best-path alignments)."""
    ms, x = synth.make_config("cfg1")   # 4501 x 16 x 45, 1000 frames
    rng = np.random.default_rng(11)
    fo = np.array([0, 250, 480, 480, 1000], np.uint32)
    segs = [linear_hmm(rng.integers(0, ms.M, n), loop=3.0, forward=0.0, skip=30.0) for n in (60, 75, 1, 140)]
    au = AlignmentAutomata.from_arc_lists(segs)
    for kernel in (capi.KERNEL_DIRECT, capi.KERNEL_AUTO):
        t = oracle_float_tables(ms) if kernel == capi.KERNEL_DIRECT else tables.float_tables(ms)
        eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX,
                         kernel=kernel)
        em, st, sc = eng.score_align_host(fo, x, au)
        s, _ = eng.score_host(x, want_best=False)
        emo, sto, sco = oracle.viterbi_align(fo, s, au)
        assert np.array_equal(em, emo) and np.array_equal(st, sto) and np.array_equal(bits(sc), bits(sco))
        so, _ = oracle.score_max(ms, x)
        emr, str_, scr = oracle.viterbi_align(fo, so, au)
        assert np.array_equal(em, emr) and np.array_equal(st, str_)
        np.testing.assert_allclose(sc, scr, rtol=1e-4)


@pytest.mark.parametrize("subset", ["0", "1"])
def test_label_entry_points_keep_tied_labels_apart(eng, oracle, monkeypatch, subset):
    """gmm_score_align_labels_host / gmm_score_baum_welch_labels_host: arcs carry labels (allophone states) of which
    several are tied to one mixture; the path / the posterior columns name the LABEL, as the reference's alignment
    items do (tests/golden/aln_tied.npz pins that behaviour of the oracle against the reference's own aligner).
    Oracle: the restated aligner over this engine's scores with one column per label (a label's score is its
    mixture's, Aligner.cc:225).  Over the compact and the full score block."""
    from rasr_b200.align import label_slots
    monkeypatch.setenv("GMM_ALIGN_SUBSET", subset)
    ms = synth.make_model(M=300, K=8, D=20, seed=33)
    rng = np.random.default_rng(33)
    n_labels = 900
    tie = rng.integers(0, ms.M, n_labels).astype(np.uint32)
    tie[:40] = tie[40:80]                                            # labels k and k+40 share a mixture
    lens = [120, 0, 35, 1, 260, 64]
    fo = np.cumsum([0] + lens).astype(np.uint32)
    x = synth.make_features(ms, int(fo[-1]), seed=34)
    segs = []
    for n in lens:
        labs = rng.integers(0, 80, max(1, n // 3))                  # mostly from the tied range
        states, ini, fin = linear_hmm(labs, loop=3.0, forward=0.0, skip=30.0)
        for s_, arcs in enumerate(states):                          # a parallel arc under the twin label, same weight
            for (tg, lab, w) in list(arcs):
                if lab < 80 and rng.random() < 0.5:
                    arcs.append((tg, (lab + 40) % 80, w))
        segs.append((states, ini, fin))
    au = AlignmentAutomata.from_arc_lists(segs)                      # arc_emission = label id
    t = tables.float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX,
                     kernel=capi.KERNEL_TENSOR)
    s, _ = eng.score_host(x, want_best=False)
    per_label = np.ascontiguousarray(s[:, tie])
    emo, sto, sco = oracle.viterbi_align(fo, per_label, au)
    slots, lo, ids, mix = label_slots(au, tie)
    lab, st, sc = eng.score_align_labels_host(fo, x, slots, lo, mix)
    assert np.array_equal(bits(sc), bits(sco)) and np.array_equal(st, sto)
    for i in range(len(lens)):
        rows = slice(int(fo[i]), int(fo[i + 1]))
        got = lab[rows]
        want = emo[rows]
        if sco[i] == FLT_MAX:
            assert np.all(got == 0xFFFFFFFF)
        else:
            assert np.array_equal(ids[lo[i]:lo[i + 1]][got], want)   # slot -> label id
    assert len(set(emo[emo < 80].tolist())) > 20                     # the tied range is really used

    wo, lists, to = oracle.baum_welch_align(fo, per_label, au)
    w, total = eng.score_baum_welch_labels_host(fo, x, slots, lo, mix)
    for i in range(len(lens)):
        rows = slice(int(fo[i]), int(fo[i + 1]))
        assert np.array_equal(lists[i], ids[lo[i]:lo[i + 1]])       # same first-use order: columns line up
        n = len(lists[i])
        if to[i] == FLT_MAX:
            assert total[i] == FLT_MAX and not w[rows, :n].any()
            continue
        assert abs(total[i] - to[i]) <= 1e-5 * max(1.0, abs(to[i]))
        assert np.max(np.abs(w[rows, :n] - wo[i]), initial=0.0) <= 1e-4


def test_label_entry_points_reject_bad_lists(eng):
    ms = synth.make_model(M=20, K=2, D=8, seed=35)
    t = tables.float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX)
    au = AlignmentAutomata.from_arc_lists([linear_hmm([0, 1, 2])])
    x = np.zeros((4, 8), np.float32)
    with pytest.raises(capi.GmmError):                               # slot 2 outside a list of two labels
        eng.score_align_labels_host([0, 4], x, au, [0, 2], [3, 4])
    with pytest.raises(capi.GmmError):                               # label tied to a mixture the model lacks
        eng.score_align_labels_host([0, 4], x, au, [0, 3], [3, 4, 20])


@pytest.mark.parametrize("subset", ["0", "1"])
def test_consumer_score_scale_is_the_scaling_wrapper(eng, oracle, monkeypatch, subset):
    """gmm_set_consumer_score_scale: the aligner sees fl(scale * score), as behind Mm::FeatureScorerScaling
    (ScaledFeatureScorer.hh:62-64) -- Viterbi bit-exact, Baum-Welch within its tolerance, on both block layouts."""
    monkeypatch.setenv("GMM_ALIGN_SUBSET", subset)
    ms = synth.make_model(M=400, K=8, D=24, seed=41)
    rng = np.random.default_rng(41)
    lens = [90, 33, 0, 150]
    fo = np.cumsum([0] + lens).astype(np.uint32)
    x = synth.make_features(ms, int(fo[-1]), seed=42)
    au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, ms.M, max(1, n // 3)), loop=3.0, forward=0.0,
                                                      skip=30.0) for n in lens])
    t = tables.float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=capi.MODE_MAX,
                     kernel=capi.KERNEL_TENSOR)
    s, _ = eng.score_host(x, want_best=False)
    scaled = (np.float32(0.75) * s).astype(np.float32)
    eng.set_consumer_score_scale(0.75)
    try:
        em, st, sc = eng.score_align_host(fo, x, au)
        w, total = eng.score_baum_welch_host(fo, x, au)
    finally:
        eng.set_consumer_score_scale(1.0)
    emo, sto, sco = oracle.viterbi_align(fo, scaled, au)
    assert np.array_equal(em, emo) and np.array_equal(st, sto) and np.array_equal(bits(sc), bits(sco))
    wo, lists, to = oracle.baum_welch_align(fo, scaled, au)
    for i in range(len(lens)):
        rows = slice(int(fo[i]), int(fo[i + 1]))
        assert np.max(np.abs(w[rows, :len(lists[i])] - wo[i]), initial=0.0) <= 1e-4
        assert abs(total[i] - to[i]) <= 1e-5 * max(1.0, abs(to[i]))
    em1, _, sc1 = eng.score_align_host(fo, x, au)                    # back at scale 1
    assert np.array_equal(bits(sc1), bits(oracle.viterbi_align(fo, s, au)[2]))
    with pytest.raises(capi.GmmError):
        eng.set_consumer_score_scale(0.0)
