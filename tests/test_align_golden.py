"""The alignment consumers against the reference's own Search::Aligner (tests/golden/aln_*.npz, written by
oracle/_ref/ref_aligner = the unmodified src/Search/Aligner.cc with its Fsa library and Speech::Alignment; generator
tests/golden/make_align_golden.py).

CPU: the oracle restatements orc_viterbi_align / orc_baum_welch_align -- this is their pin.
GPU: gmm_align_device and gmm_baum_welch_device directly against the reference's answers.

Viterbi is bit-exact: the same emission at every frame (exact f32 ties included, aln_ties) and the same f32 path
score.  Baum-Welch: the reference keeps arc posteriors as f32 -log values and combines them in the f32 log semiring
(Aligner.cc:697-709), the oracle works in f64 and the device in f32 potentials: state posteriors agree to 1e-5 (oracle)
/ 1e-4 (device) absolute, totals to 1e-6 / 1e-5 relative."""
import numpy as np
import pytest

from conftest import align_golden_names, load_align_golden

FLT_MAX = np.finfo(np.float32).max
NAMES = align_golden_names()


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def dense_reference_posteriors(z, i, emission_list, frames):
    """The reference's items of segment i as a [frames x len(list)] matrix (an item per (time, emission))."""
    col = {int(e): j for j, e in enumerate(emission_list)}
    out = np.zeros((frames, max(1, len(col))), np.float64)
    a, b = int(z["bw_item_offsets"][i]), int(z["bw_item_offsets"][i + 1])
    seen = set()
    for t, e, w in zip(z["bw_time"][a:b], z["bw_emission"][a:b], z["bw_weight"][a:b]):
        assert (int(t), int(e)) not in seen, "combineItems leaves one item per (time, emission)"
        seen.add((int(t), int(e)))
        out[int(t), col[int(e)]] = w
    return out[:, :len(col)]


def label_block(scores, z):
    """aln_tied: arcs carry labels the acoustic model ties to the columns of `scores` -- a label's score is its
    emission's (Aligner.cc:225), so the block with one column per label is scores[:, emission_of_label]."""
    return np.ascontiguousarray(scores[:, z["emission_of_label"]]) if "emission_of_label" in z.files else scores


def test_fixtures_present():
    assert len(NAMES) >= 6 and "aln_tied" in NAMES


@pytest.mark.parametrize("name", NAMES)
def test_oracle_viterbi_is_the_reference_aligner(oracle, name):
    fo, scores, au, z = load_align_golden(name)
    scores = label_block(scores, z)
    em, st, sc = oracle.viterbi_align(fo, scores, au)
    assert np.array_equal(bits(sc), bits(z["vit_score"]))
    assert np.array_equal(em, z["vit_emission"])
    # states of the path: every step is an arc of the automaton that emits what the reference aligned
    for i in range(len(fo) - 1):
        if sc[i] == FLT_MAX:
            continue
        s0 = int(au.state_offsets[i])
        s = int(au.initial_state[i])
        for t in range(int(fo[i]), int(fo[i + 1])):
            a0, a1 = int(au.arc_offsets[s0 + s]), int(au.arc_offsets[s0 + s + 1])
            assert any(au.arc_target[a] == st[t] and au.arc_emission[a] == em[t] for a in range(a0, a1))
            s = int(st[t])
        assert np.isfinite(au.final_weight[s0 + s])


@pytest.mark.parametrize("name", NAMES)
def test_oracle_baum_welch_is_the_reference_aligner(oracle, name):
    fo, scores, au, z = load_align_golden(name)
    scores = label_block(scores, z)
    w, lists, total = oracle.baum_welch_align(fo, scores, au)
    for i in range(len(fo) - 1):
        frames = int(fo[i + 1]) - int(fo[i])
        if z["bw_score"][i] == FLT_MAX:
            assert total[i] == FLT_MAX and not w[i].any() and z["bw_item_offsets"][i] == z["bw_item_offsets"][i + 1]
            continue
        assert abs(total[i] - -float(z["bw_total_inv"][i])) <= 1e-6 * max(1.0, abs(total[i]))
        assert abs(total[i] - float(z["bw_score"][i])) <= 1e-5 * max(1.0, abs(total[i]))
        want = dense_reference_posteriors(z, i, lists[i], frames)
        assert np.max(np.abs(w[i] - want), initial=0.0) <= 1e-5
        if frames:
            np.testing.assert_allclose(want.sum(axis=1), 1.0, atol=1e-9)   # normalizeWeights, per frame


# ----------------------------------------------------------------------------------------------------------------
# the device consumers, directly against the reference's answers


@pytest.fixture(scope="module")
def eng():
    from rasr_b200 import capi
    e = capi.Engine(0)
    yield e
    e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["ranked", "generic"])
@pytest.mark.parametrize("name", NAMES)
def test_device_viterbi_is_the_reference_aligner(eng, monkeypatch, name, kernel):
    import torch
    if kernel == "generic":
        monkeypatch.setenv("GMM_ALIGN_KERNEL", "generic")
    else:
        monkeypatch.delenv("GMM_ALIGN_KERNEL", raising=False)
    fo, scores, au, z = load_align_golden(name)
    scores = label_block(scores, z)
    em, st, sc = eng.align_device(fo, torch.from_numpy(scores).cuda(), au, ld=scores.shape[1])
    assert np.array_equal(bits(sc), bits(z["vit_score"]))
    assert np.array_equal(em, z["vit_emission"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_device_baum_welch_is_the_reference_aligner(eng, name):
    import torch
    fo, scores, au, z = load_align_golden(name)
    scores = label_block(scores, z)
    lo, lst = eng.emission_lists(au, scores.shape[1])
    ld_w = max(1, int(np.max(np.diff(lo)))) + 1
    w = torch.full((scores.shape[0], ld_w), -1.0, device="cuda")
    total = eng.baum_welch_device(fo, torch.from_numpy(scores).cuda(), au, w)
    w = w.cpu().numpy()
    for i in range(len(fo) - 1):
        rows = slice(int(fo[i]), int(fo[i + 1]))
        lst_i = lst[lo[i]:lo[i + 1]]
        if z["bw_score"][i] == FLT_MAX:
            assert total[i] == FLT_MAX and not w[rows, :len(lst_i)].any()
            continue
        assert abs(total[i] - -float(z["bw_total_inv"][i])) <= 1e-5 * max(1.0, abs(float(total[i])))
        want = dense_reference_posteriors(z, i, lst_i, rows.stop - rows.start)
        assert np.max(np.abs(w[rows, :len(lst_i)] - want), initial=0.0) <= 1e-4
