"""CPU: host-side logic -- table packing against the oracle, the buffered-scorer state machine (with the GPU
engine replaced by a recording stub), the GMS container, and the C-ABI library's exported symbols."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden_names, load_golden
from rasr_b200 import capi, feature_scorer, synth, tables
from rasr_b200.gms import MixtureSetData


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("kw", [dict(M=13, K=16, D=45), dict(M=7, K=5, D=9, covariance="density", ragged=True,
                                                             shared=True, zero_weight_frac=0.25)])
def test_tables_match_oracle_exact_mode(oracle, kw):
    ms = synth.make_model(**kw)
    t = tables.float_tables(ms, mixture_weight_scale=0.7, gaussian_scale=1.3)
    isr, logn, w2 = oracle.float_tables(ms, mixture_weight_scale=0.7, gaussian_scale=1.3, exact_rsqrt=True)
    assert np.array_equal(bits(t["isr"]), bits(isr))
    assert np.array_equal(bits(t["log_norm"]), bits(logn))
    assert np.array_equal(bits(t["w2"]), bits(w2))
    q = tables.quant_tables(ms)
    o = oracle.simd_prepare(ms, exact_rsqrt=True)
    assert np.array_equal(q["qmean"], o["qmean"])
    assert np.array_equal(q["qconst"], o["qconst"])
    assert np.array_equal(q["cov_idx"], o["qcov"])
    assert np.array_equal(bits(q["isr_scaled"]), bits(o["isrs"]))
    assert np.float32(q["scaling_squared"]) == np.float32(o["scaling_squared"])


def test_quantize_u8_matches_oracle(oracle):
    rng = np.random.default_rng(3)
    v = np.concatenate([rng.standard_normal(1000) * 60, [0.5, -0.5, 1.5, 2.5, -2.5, 126.5, 127.5, -128.5, 1e20, -1e20,
                                                          3e9, -3e9, np.inf, -np.inf]]).astype(np.float32)
    ones = np.ones((1, v.shape[0]), np.float32)
    qx = oracle.simd_quantize_features(v.reshape(1, -1), ones)[0, 0, : v.shape[0]]
    assert np.array_equal(tables.quantize_u8(v), qx)


def test_gms_roundtrip(tmp_path):
    ms = synth.make_model(M=5, K=3, D=7, covariance="mixture", shared=True)
    p = str(tmp_path / "m.gms")
    ms.save(p)
    ms2 = MixtureSetData.load(p)
    for f in ("mix_offsets", "mix_density", "mix_logw", "dens_mean", "dens_cov", "means", "variances"):
        assert np.array_equal(getattr(ms, f), getattr(ms2, f))
    with pytest.raises(ValueError):
        MixtureSetData(2, [0, 1], [0], [0.0], [0], [0], [[1, 2]], [[1.0, -1.0]])


def test_capi_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "gmm_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(gmm_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == {s[0] for s in capi.SYMBOLS}
    lib = ctypes.CDLL(capi.LIB_PATH)  # must load without a GPU
    for name in declared:
        assert hasattr(lib, name), name
    capi.load()
    assert b"sm_100a" in capi.load().gmm_version()


def test_engine_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.GmmError):
        capi.Engine(0)


class _StubEngine:
    """Stands in for capi.Engine: score of (frame t, mixture m) = x[t,0] + m, so the frame identity is visible."""

    def __init__(self, device=0):
        self.batches = []

    def upload_float(self, offsets, mean, isr, cov_idx, w2, log_norm, mode=0, kernel=0):
        self.M = offsets.shape[0] - 1

    def score_host(self, x, want_best=True, **kw):
        self.batches.append(x.shape[0])
        s = x[:, :1] + np.arange(self.M, dtype=np.float32)[None, :]
        return s.astype(np.float32), (np.zeros(s.shape, np.uint32) if want_best else None)


@pytest.fixture
def stub(monkeypatch):
    monkeypatch.setattr(capi, "Engine", _StubEngine)
    return _StubEngine


def _frames(n, D):
    x = np.zeros((n, D), np.float32)
    x[:, 0] = np.arange(n) * 100
    return x


def test_streaming_buffer_protocol(stub):
    """The caller loop of Speech/Recognizer.cc:272-282,198-206 against a buffer-size-4 scorer."""
    ms = synth.make_model(M=3, K=2, D=5)
    fs = feature_scorer.GpuFeatureScorer(ms, buffer_size=4)
    assert fs.isBuffered() and fs.bufferSize() == 4 and fs.bufferEmpty()
    got = []
    for f in _frames(10, 5):
        if fs.isBuffered() and not fs.bufferFilled():
            fs.addFeature(f)
        else:
            got.append(fs.getScorer(f))
            got[-1].score(0)  # a decoder consumes the frame's scores before feeding the next one
    while not fs.bufferEmpty():
        got.append(fs.flush())
    assert len(got) == 10
    assert [s.score(1) for s in got] == [t * 100 + 1 for t in range(10)]  # FIFO order, one scorer per frame
    assert got[0].nEmissions() == 3
    assert sum(fs.engine.batches) == 10 and max(fs.engine.batches) <= 4
    with pytest.raises(AssertionError):
        fs.flush()
    fs.addFeature(np.zeros(5, np.float32)), fs.addFeature(np.zeros(5, np.float32)), fs.addFeature(np.zeros(5, np.float32))
    assert fs.bufferFilled()
    with pytest.raises(AssertionError):
        fs.addFeature(np.zeros(5, np.float32))
    fs.reset()
    assert fs.bufferEmpty() and not fs.bufferFilled()


def test_segment_mode_scores_in_one_batch(stub):
    """buffer-size 0: never filled (Nn/TrainerFeatureScorer.cc:220-230); AlignmentNode.cc:283-306 loop."""
    ms = synth.make_model(M=3, K=2, D=5)
    fs = feature_scorer.createFeatureScorer({"feature-scorer-type": "diagonal-maximum", "buffer-size": 0}, ms)
    fs.reset()
    scorers = []
    for f in _frames(37, 5):
        if fs.isBuffered() and not fs.bufferFilled():
            fs.addFeature(f)
        else:
            scorers.append(fs.getScorer(f))
    while not fs.bufferEmpty():
        scorers.append(fs.flush())
    assert fs.bufferSize() == 0xFFFFFFFF and len(scorers) == 37
    assert scorers[36].score(2) == 3602 and scorers[0].score(0) == 0
    assert fs.engine.batches == [37]
    # time-indexed access in segment mode (FeatureScorer.hh:157-163, SegmentwiseAlignmentGenerator.cc:90-156)
    assert fs.hasTimeIndexedCache() and fs.getTimeIndexedScorer(36) is scorers[36]
    assert fs.getTimeIndexedScorer(4).score(1) == scorers[4].score(1)
    buffered = feature_scorer.createFeatureScorer({"feature-scorer-type": "diagonal-maximum", "buffer-size": 8}, ms)
    assert not buffered.hasTimeIndexedCache()
    with pytest.raises(RuntimeError, match="time-indexed scores not available"):
        buffered.getTimeIndexedScorer(0)
    # contexts outlive the next segment (the aligner keeps them)
    fs.reset()
    s = fs.getScorer(_frames(1, 5)[0] + 7)
    assert s.score(0) == 7 and scorers[5].score(0) == 500


def test_unknown_scorer_type_and_dimension(stub):
    ms = synth.make_model(M=3, K=2, D=5)
    with pytest.raises(ValueError):
        feature_scorer.GpuFeatureScorer(ms, "no-such-scorer")
    fs = feature_scorer.GpuFeatureScorer(ms)
    with pytest.raises(AssertionError):
        fs.addFeature(np.zeros(4, np.float32))


@pytest.mark.parametrize("name", golden_names()[:2])
def test_golden_fixture_loads(name):
    ms, z = load_golden(name)
    assert z["max_score"].shape == (z["x"].shape[0], ms.M)


def test_label_slots_keep_tied_labels_apart():
    """rasr_b200.align.label_slots: per segment the distinct arc labels in order of first use become slots; labels tied
    to one mixture stay apart (the reference's alignment items name the label, Aligner.cc:185, 205)."""
    from rasr_b200.align import AlignmentAutomata, label_slots
    tie = np.array([5, 5, 7, 2, 7, 9], np.uint32)                       # labels 0,1 -> mixture 5; 2,4 -> 7
    seg_a = ([[(1, 4, 0.0), (1, 2, 0.5), (0, 4, 1.0)], [(1, 1, 0.0)]], 0, {1: 0.0})
    seg_b = ([[(0, 3, 0.0)], []], 0, {0: 1.0})
    seg_c = ([[]], 0, {0: 0.0})                                         # no arcs: an empty label list
    au = AlignmentAutomata.from_arc_lists([seg_a, seg_b, seg_c])
    slots, lo, ids, mix = label_slots(au, tie)
    assert lo.tolist() == [0, 3, 4, 4]
    assert ids.tolist() == [4, 2, 1, 3] and mix.tolist() == [7, 7, 5, 2]
    assert slots.arc_emission.tolist() == [0, 1, 0, 2, 0]
    assert np.array_equal(slots.arc_target, au.arc_target) and np.array_equal(slots.arc_weight, au.arc_weight)
    for i in range(3):                                                  # slot -> label -> the original arc labels
        a0, a1 = int(au.arc_offsets[au.state_offsets[i]]), int(au.arc_offsets[au.state_offsets[i + 1]])
        assert np.array_equal(ids[lo[i]:lo[i + 1]][slots.arc_emission[a0:a1]], au.arc_emission[a0:a1])


def test_aligner_mirror_configuration_errors():
    """rasr_b200.aligner.Aligner refuses what the device consumers do not do (as rasr/GpuAligner.cc: criticalError)."""
    from rasr_b200.aligner import Aligner
    with pytest.raises(ValueError):
        Aligner({"min-weight": 0.01})
    with pytest.raises(ValueError):
        Aligner({"mode": "forward-backward"})
    a = Aligner({"mode": "baum-welch"})
    assert a.getMode() == Aligner.modeBaumWelch and not a.reachedFinalState()
    a.selectMode(Aligner.modeViterbi)
    assert a.getMode() == Aligner.modeViterbi
    from rasr_b200.align import linear_hmm
    a.setModel(linear_hmm([3, 4]), np.arange(10))
    a.feed([])                                                          # no frames: the start state is not final
    assert not a.reachedFinalState() and a.getAlignment() == []
    a.setModel(([[(0, 1, 0.0)]], 0, {0: 2.5}), np.arange(10))
    a.feed([])                                                          # no frames, final start state: its weight
    assert a.reachedFinalState() and a.alignmentScore() == 2.5
