"""GPU: subset scoring (gmm_score_subset_device) gives, for the listed mixtures of every segment, the very bits the
full tensor-kernel pass gives, and stays within the float bar of the oracle."""
import numpy as np
import pytest
import torch

from rasr_b200 import capi, synth, tables

pytestmark = pytest.mark.gpu


@pytest.fixture()
def eng():
    e = capi.Engine(0)
    yield e
    e.close()


def upload(eng, ms, mode=capi.MODE_MAX):
    t = tables.float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=mode,
                     kernel=capi.KERNEL_TENSOR)


@pytest.mark.parametrize("M,K,D,cov,mode", [(4501, 16, 45, "pooled", capi.MODE_MAX), (300, 8, 20, "density", capi.MODE_MAX),
                                            (97, 32, 33, "pooled", capi.MODE_MAX), (200, 16, 24, "pooled", capi.MODE_SUM),
                                            (64, 3, 9, "mixture", capi.MODE_MAX)])
def test_subset_equals_full_scoring(eng, oracle, M, K, D, cov, mode):
    ms = synth.make_model(M=M, K=K, D=D, seed=M + K, covariance=cov)
    rng = np.random.default_rng(M)
    lens = [1, 127, 128, 129, 0, 700, 33] + [int(n) for n in rng.integers(1, 400, 12)]
    fo = np.cumsum([0] + lens).astype(np.uint32)
    x = synth.make_features(ms, int(fo[-1]), seed=3)
    upload(eng, ms, mode)
    dx = torch.from_numpy(x).cuda()
    full = torch.empty((x.shape[0], M), device="cuda")
    eng.score_device(x.shape[0], dx, full)
    lists = [rng.choice(M, size=int(n), replace=False) for n in rng.integers(1, min(M, 350), len(lens))]
    lists[3] = np.arange(M)[::-1].copy()            # a whole model, backwards
    lists[5] = np.array([M - 1])                    # a single mixture
    lo = np.cumsum([0] + [len(v) for v in lists]).astype(np.uint32)
    ld = eng.subset_ld(max(len(v) for v in lists))
    sub = torch.full((x.shape[0], ld), -1.0, device="cuda")
    eng.score_subset_device(fo, lo, np.concatenate(lists), dx, sub, ld)
    eng.sync()
    full, sub = full.cpu().numpy(), sub.cpu().numpy()
    for s, v in enumerate(lists):
        rows = slice(int(fo[s]), int(fo[s + 1]))
        assert np.array_equal(sub[rows, :len(v)].view(np.uint32), full[rows][:, v].view(np.uint32)), s
    if mode == capi.MODE_MAX:
        ref, _ = oracle.score_max(ms, x[:int(fo[2])])
        rows = slice(int(fo[1]), int(fo[2]))
        r = ref[rows][:, lists[1]]
        g = sub[rows, :len(lists[1])]
        assert np.max(np.abs(g - r) / np.maximum(np.abs(r), 1)) <= 1e-4


def test_subset_needs_a_regular_model(eng):
    ms = synth.make_model(M=40, K=6, D=10, ragged=True, seed=2)
    upload(eng, ms)
    dx = torch.zeros((4, 10), device="cuda")
    out = torch.zeros((4, 64), device="cuda")
    with pytest.raises(capi.GmmError, match="same number"):
        eng.score_subset_device([0, 4], [0, 2], [1, 2], dx, out, 64)
    ms = synth.make_model(M=40, K=8, D=10, seed=2)
    upload(eng, ms)
    with pytest.raises(capi.GmmError, match="narrower"):
        eng.score_subset_device([0, 4], [0, 40], np.arange(40), dx, out, 8)
    with pytest.raises(capi.GmmError, match="out of range"):
        eng.score_subset_device([0, 4], [0, 1], [40], dx, out, 64)


@pytest.mark.parametrize("align_kernel", ["ranked", "generic"])
def test_alignment_over_compact_score_blocks(eng, oracle, monkeypatch, align_kernel):
    """gmm_score_align_host scores only the emissions each automaton can ask for (compact block, one column per
    entry of the segment's emission list): paths, states and path scores must be the bits of the full-block run, and
    the oracle's Viterbi path over the engine's own scores."""
    from rasr_b200.align import AlignmentAutomata, linear_hmm
    if align_kernel == "generic":
        monkeypatch.setenv("GMM_ALIGN_KERNEL", "generic")
    ms = synth.make_model(M=1200, K=16, D=45, seed=5)
    rng = np.random.default_rng(3)
    lens = [300, 0, 129, 1, 700, 64]
    fo = np.cumsum([0] + lens).astype(np.uint32)
    x = synth.make_features(ms, int(fo[-1]), seed=9)
    segs = [linear_hmm(rng.integers(0, ms.M, n), loop=3.0, forward=0.0, skip=30.0) for n in (40, 5, 33, 1, 290, 17)]
    au = AlignmentAutomata.from_arc_lists(segs)
    upload(eng, ms)
    out = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("GMM_ALIGN_SUBSET", flag)
        n0 = eng.launch_count
        out[flag] = eng.score_align_host(fo, x, au)
        out[flag + "n"] = eng.launch_count - n0
    for a, b in zip(out["0"], out["1"]):
        assert np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))
    assert out["1n"] == out["0n"] + 2            # the compact path adds the tile gather and its table copy
    monkeypatch.delenv("GMM_ALIGN_SUBSET")
    auto = eng.score_align_host(fo, x, au)       # the work estimate picks the compact path here by itself
    assert np.array_equal(auto[0], out["1"][0])
    s, _ = eng.score_host(x, want_best=False)
    emo, sto, sco = oracle.viterbi_align(fo, s, au)
    assert np.array_equal(out["1"][0], emo) and np.array_equal(out["1"][1], sto)
    assert np.array_equal(out["1"][2].view(np.uint32), sco.view(np.uint32))


@pytest.mark.parametrize("subset", ["0", "1"])
def test_alignment_pipeline_across_two_lanes(oracle, monkeypatch, subset):
    """Groups of segments alternate between the handle's two lanes (stream + scratch + pinned staging each): with
    groups of 2 segments a 9-segment call runs 5 groups, reuses each lane twice, and must give the paths of the
    one-group call and of the oracle; the training pass must accumulate every aligned frame exactly once."""
    from rasr_b200.align import AlignmentAutomata, linear_hmm
    monkeypatch.setenv("GMM_ALIGN_SUBSET", subset)
    monkeypatch.setenv("GMM_HOST_THREADS", "3")
    e = capi.Engine(0)
    try:
        ms = synth.make_model(M=700, K=16, D=45, seed=8)
        rng = np.random.default_rng(5)
        lens = [200, 129, 0, 64, 333, 1, 128, 90, 257]
        fo = np.cumsum([0] + lens).astype(np.uint32)
        x = synth.make_features(ms, int(fo[-1]), seed=4)
        segs = [linear_hmm(rng.integers(0, ms.M, n), loop=3.0, forward=0.0, skip=30.0)
                for n in (50, 30, 4, 20, 120, 1, 60, 25, 100)]
        au = AlignmentAutomata.from_arc_lists(segs)
        upload(e, ms)
        monkeypatch.setenv("GMM_ALIGN_GROUP", "1000")
        one = e.score_align_host(fo, x, au)
        monkeypatch.setenv("GMM_ALIGN_GROUP", "2")
        two = e.score_align_host(fo, x, au)
        for a, b in zip(one, two):
            assert np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))
        s, _ = e.score_host(x, want_best=False)
        emo, sto, sco = oracle.viterbi_align(fo, s, au)
        assert np.array_equal(two[0], emo) and np.array_equal(two[1], sto)
        assert np.array_equal(two[2].view(np.uint32), sco.view(np.uint32))
        # training pass over the lanes: unweighted counts are exact, one per aligned frame
        e.create_accumulators(tables.mean_index(ms), ms.n_means)
        e.score_align_host(fo, x, au, accumulate=True, want_paths=False)
        acc = e.read_accumulators(ms.variances.shape[0])
        aligned = int(np.sum(emo != 0xFFFFFFFF))
        assert int(round(float(np.sum(acc["mix_weight"])))) == aligned
    finally:
        e.close()
