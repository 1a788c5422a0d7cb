"""GPU: rasr_b200.aligner.Aligner -- Search::Aligner's calls (setModel / feed / reachedFinalState / alignmentScore /
getAlignment) over the device consumers, the Python statement of rasr/GpuAligner.cc -- against the oracle's restatement
of the reference aligner (pinned against the reference's own Search::Aligner, tests/test_align_golden.py) run on this
engine's scores.  The C++ class itself is checked in process against the reference's aligner (test_gpu_rasr_plugin.py)."""
import numpy as np
import pytest

from rasr_b200 import synth
from rasr_b200.align import AlignmentAutomata, linear_hmm
from rasr_b200.aligner import Aligner
from rasr_b200.feature_scorer import GpuFeatureScorer

pytestmark = pytest.mark.gpu
FLT_MAX = float(np.finfo(np.float32).max)


def segment_scorers(fs, x):
    """the caller loop of Speech/AlignmentNode.cc:283-306"""
    scorers = []
    fs.reset()
    for f in x:
        if fs.isBuffered() and not fs.bufferFilled():
            fs.addFeature(f)
        else:
            scorers.append(fs.getScorer(f))
    while not fs.bufferEmpty():
        scorers.append(fs.flush())
    return scorers


@pytest.mark.parametrize("mode", [Aligner.modeViterbi, Aligner.modeBaumWelch])
def test_aligner_mirror_matches_the_restated_reference_aligner(oracle, mode):
    ms = synth.make_model(M=120, K=8, D=20, seed=51)
    rng = np.random.default_rng(51)
    tie = np.concatenate([np.arange(ms.M), np.arange(ms.M)]).astype(np.uint32)     # labels l and l + M share mixture l
    fs = GpuFeatureScorer(ms, "diagonal-maximum", buffer_size=0, assigning=False)
    aligner = Aligner({"mode": mode})
    assert aligner.getMode() == mode
    scale = 0.75
    for T in (1, 40, 0, 170):
        labels = rng.integers(0, 2 * ms.M, max(1, T // 3))
        if len(labels) > 3:
            labels[2] = (labels[1] + ms.M) % (2 * ms.M)                          # a sibling tied to the same mixture
        model = linear_hmm(labels, loop=3.0, forward=0.0, skip=30.0)
        x = synth.make_features(ms, T, seed=52 + T) if T else np.zeros((0, ms.D), np.float32)
        aligner.setModel(model, tie)
        scorers = segment_scorers(fs, x)
        aligner.feed(scorers, scale=scale)
        assert fs.n_batches == 0                                                   # nothing was scored on the side
        au = AlignmentAutomata.from_arc_lists([model])
        s = fs.engine.score_host(x, want_best=False)[0] if T else np.zeros((0, ms.M), np.float32)
        per_label = np.ascontiguousarray((np.float32(scale) * s)[:, tie]).astype(np.float32)
        if mode == Aligner.modeViterbi:
            emo, _, sco = oracle.viterbi_align([0, T], per_label, au)
            assert np.float32(aligner.alignmentScore()).view(np.uint32) == sco[0].view(np.uint32)
            assert aligner.reachedFinalState() == (sco[0] != np.float32(FLT_MAX))
            items = aligner.getAlignment()
            if aligner.reachedFinalState():
                assert [(t, l) for t, l, _ in items] == list(enumerate(emo.tolist())) and all(w == 1.0 for *_, w in items)
            else:
                assert items == []
        else:
            wo, lists, to = oracle.baum_welch_align([0, T], per_label, au)
            assert aligner.reachedFinalState() == (to[0] != FLT_MAX)
            if not aligner.reachedFinalState():
                assert aligner.getAlignment() == []
                continue
            assert abs(aligner.alignmentScore() - to[0]) <= 1e-5 * max(1.0, abs(to[0]))
            dense = np.zeros((T, len(lists[0])))
            col = {int(l): j for j, l in enumerate(lists[0])}
            last = (-1, 2.0)
            for t, l, w in aligner.getAlignment():
                dense[t, col[l]] += w
                assert (t, -w) >= (last[0], -last[1]) and w > 0                   # by time, then decreasing weight
                last = (t, w)
            assert np.max(np.abs(dense - wo[0]), initial=0.0) <= 1e-4


def test_aligner_mirror_refuses_what_it_does_not_do():
    with pytest.raises(ValueError):
        Aligner({"min-weight": 0.01})
    with pytest.raises(ValueError):
        Aligner({"mode": "forward-backward"})
    ms = synth.make_model(M=20, K=2, D=8, seed=53)
    fs = GpuFeatureScorer(ms, "SIMD-diagonal-maximum", buffer_size=0)
    aligner = Aligner()
    aligner.setModel(linear_hmm([1, 2]), np.arange(ms.M))
    with pytest.raises(ValueError):
        aligner.feed(segment_scorers(fs, synth.make_features(ms, 4)))
