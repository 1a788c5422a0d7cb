"""GPU parity of the front end fused into the frame upload (gmm_set_lda / gmm_lda_device):
bit-exact against the oracle and the golden outputs of the reference's own matrix-vector product, and every *_host
entry point fed with base features gives what it gives for the transformed features."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR, oracle_float_tables, oracle_quant_tables
from rasr_b200 import capi, synth, tables
from rasr_b200.align import AlignmentAutomata, linear_hmm

pytestmark = pytest.mark.gpu


@pytest.fixture()
def eng():
    e = capi.Engine(0)
    yield e
    e.close()


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def device_lda(eng, base, fo, D):
    d_base = torch.from_numpy(base).cuda()
    d_x = torch.empty((base.shape[0], D), device="cuda")
    eng.lda_device(fo, d_base, d_x)
    eng.sync()
    return d_x.cpu().numpy()


def test_matrix_product_matches_reference_golden(eng):
    z = np.load(os.path.join(GOLDEN_DIR, "lda_matmul.npz"))
    for name in ("lda45x144", "odd7x5", "wide3x1000"):
        m, v, want = z[name + "_matrix"], z[name + "_in"], z[name + "_out"]
        eng.set_lda(m, base_dim=m.shape[1], length=1, right=0)      # window of one frame: the product alone
        got = device_lda(eng, v, [0, v.shape[0]], m.shape[0])
        assert np.array_equal(bits(got), bits(want)), name


@pytest.mark.parametrize("name", ["lda16x9", "causal7x3", "lookahead5x6", "wide3x17"])
def test_window_and_product_match_reference_golden(eng, name):
    """tests/golden/lda_window.npz: the reference's own sliding-window node (Signal::DelayNode, wired as lda.flow)
    followed by its own matrix product -- the device front end against both, bit for bit."""
    z = np.load(os.path.join(GOLDEN_DIR, "lda_window.npz"))
    base, fo, m = z[name + "_base"], z[name + "_frame_offsets"], z[name + "_matrix"]
    length, right = int(z[name + "_length"]), int(z[name + "_right"])
    eng.set_lda(m, base.shape[1], length, right)
    assert np.array_equal(bits(device_lda(eng, base, fo, m.shape[0])), bits(z[name + "_out"]))
    eye = np.eye(base.shape[1] * length, dtype=np.float32)
    eng.set_lda(eye, base.shape[1], length, right)
    assert np.array_equal(bits(device_lda(eng, base, fo, eye.shape[0])), bits(z[name + "_window"]))


@pytest.mark.parametrize("base_dim,length,right,D", [(16, 9, 4, 45), (16, 11, 5, 45), (7, 3, 0, 5), (40, 1, 0, 40),
                                                     (33, 17, 16, 128)])
def test_windowed_lda_matches_oracle(eng, oracle, base_dim, length, right, D):
    rng = np.random.default_rng(base_dim + length)
    fo = np.cumsum([0] + [int(n) for n in rng.integers(0, 90, 40)] + [1, 2, 700]).astype(np.uint32)
    base = (rng.standard_normal((int(fo[-1]), base_dim)) * 10).astype(np.float32)
    m = rng.standard_normal((D, base_dim * length)).astype(np.float32)
    eng.set_lda(m, base_dim, length, right)
    got = device_lda(eng, base, fo, D)
    assert np.array_equal(bits(got), bits(oracle.lda(base, fo, length, right, m)))


def test_host_entry_points_take_base_features(eng, oracle):
    ms = synth.make_model(M=61, K=4, D=12, seed=14)
    rng = np.random.default_rng(14)
    fo = np.array([0, 50, 50, 131, 200], np.uint32)
    base = rng.standard_normal((200, 5)).astype(np.float32)
    m = (rng.standard_normal((12, 5 * 7)) * 0.3).astype(np.float32)
    x_seg = oracle.lda(base, fo, 7, 3, m)              # windows per segment (score_align)
    x_one = oracle.lda(base, [0, 200], 7, 3, m)        # the whole call as one segment (score, accumulate)
    t = oracle_float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], kernel=capi.KERNEL_DIRECT)
    plain, plain_best = eng.score_host(x_one)
    segs = [linear_hmm(rng.integers(0, ms.M, n)) for n in (9, 3, 20, 11)]
    au = AlignmentAutomata.from_arc_lists(segs)
    want_path = eng.score_align_host(fo, x_seg, au)
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    mixture = rng.integers(0, ms.M, 200).astype(np.uint32)
    want_best = eng.accumulate_host(x_one, mixture)
    want_acc = eng.read_accumulators(ms.n_cov)
    eng.reset_accumulators()

    eng.set_lda(m, 5, 7, 3)
    s, b = eng.score_host(base)
    assert np.array_equal(bits(s), bits(plain)) and np.array_equal(b, plain_best)
    got_path = eng.score_align_host(fo, base, au)
    assert all(np.array_equal(a, c) for a, c in zip(got_path, want_path))
    assert np.array_equal(eng.accumulate_host(base, mixture), want_best)
    got_acc = eng.read_accumulators(ms.n_cov)
    np.testing.assert_allclose(got_acc["mean_sum"], want_acc["mean_sum"], rtol=1e-12, atol=1e-12)
    with pytest.raises(ValueError):
        eng.score_host(x_one)                          # model-dimension features are no longer accepted
    eng.clear_lda()
    s2, _ = eng.score_host(x_one)
    assert np.array_equal(bits(s2), bits(plain))

    # quantized host path (frames chunked inside; the front end runs once over the whole block)
    qt, _ = oracle_quant_tables(ms)
    eng.upload_quant(qt["offsets"], qt["qmean"], qt["cov_idx"], qt["qconst"], qt["isr_scaled"], qt["scaling_squared"])
    want_q = eng.quant_score_host(x_one)
    eng.set_lda(m, 5, 7, 3)
    got_q = eng.quant_score_host(base)
    assert np.array_equal(got_q["q"], want_q["q"]) and np.array_equal(got_q["best"], want_q["best"])


def test_lda_errors(eng):
    with pytest.raises(capi.GmmError, match="no front end"):
        eng.lda_device([0, 1], torch.zeros(4, device="cuda"), torch.zeros(4, device="cuda"))
    with pytest.raises(capi.GmmError, match="right < length"):
        eng.set_lda(np.zeros((2, 6), np.float32), 2, 3, 3)
    ms = synth.make_model(M=3, K=2, D=8, seed=1)
    t = oracle_float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"])
    eng.set_lda(np.zeros((6, 4), np.float32), 4, 1, 0)   # produces 6 dimensions, the model takes 8
    with pytest.raises(capi.GmmError, match="different dimension"):
        eng.score_host(np.zeros((3, 4), np.float32))
