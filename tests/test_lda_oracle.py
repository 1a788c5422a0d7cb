"""CPU: the front end restatement (oracle/gmm_oracle.c orc_lda / orc_matrix_vector) against the golden outputs of
the reference's Math::Matrix<f32> * Math::Vector<f32> and of its sliding-window node (Signal::DelayNode in a
Flow::Network wired as share/lda.flow) and, where they are built, against those binaries themselves."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_matrix_vector_matches_reference_golden(oracle):
    z = np.load(os.path.join(GOLDEN_DIR, "lda_matmul.npz"))
    for name in ("lda45x144", "odd7x5", "wide3x1000"):
        got = oracle.matrix_vector(z[name + "_matrix"], z[name + "_in"])
        assert np.array_equal(bits(got), bits(z[name + "_out"])), name


def test_matrix_vector_matches_compiled_reference(oracle):
    if not os.path.exists(oracle.REF_MATMUL):
        pytest.skip("oracle/_ref/ref_matmul not built")
    rng = np.random.default_rng(5)
    m = rng.standard_normal((33, 77)).astype(np.float32)
    v = (rng.standard_normal((50, 77)) * 1e3).astype(np.float32)
    assert np.array_equal(bits(oracle.matrix_vector(m, v)), bits(oracle.run_ref_matmul(m, v)))


def test_window_clamps_to_the_segment(oracle):
    base = np.arange(7, dtype=np.float32).reshape(7, 1) + 1          # frames 1..7, two segments [0,4) [4,7)
    eye = np.eye(3, dtype=np.float32)                                # output = the window itself
    out = oracle.lda(base, [0, 4, 7], length=3, right=1, matrix=eye)
    want = [[1, 1, 2], [1, 2, 3], [2, 3, 4], [3, 4, 4], [5, 5, 6], [5, 6, 7], [6, 7, 7]]
    assert np.array_equal(out, np.array(want, np.float32))
    out = oracle.lda(base, [0, 4, 4, 7], length=3, right=0, matrix=eye)      # causal window, an empty segment
    assert np.array_equal(out[:4], np.array([[1, 1, 1], [1, 1, 2], [1, 2, 3], [2, 3, 4]], np.float32))
    one = oracle.lda(base[:1], [0, 1], length=5, right=2, matrix=np.ones((1, 5), np.float32))
    assert one[0, 0] == 5                                           # a one-frame segment repeats its frame


def test_lda_equals_window_then_product(oracle):
    rng = np.random.default_rng(6)
    base = rng.standard_normal((40, 16)).astype(np.float32)
    m = rng.standard_normal((45, 16 * 9)).astype(np.float32)
    fo = [0, 13, 40]
    out = oracle.lda(base, fo, length=9, right=4, matrix=m)
    for s in range(2):
        b, e = fo[s], fo[s + 1]
        win = np.stack([np.concatenate([base[min(max(t + r, b), e - 1)] for r in range(-4, 5)]) for t in range(b, e)])
        assert np.array_equal(bits(out[b:e]), bits(oracle.matrix_vector(m, win)))


WINDOW_CASES = ("lda16x9", "causal7x3", "lookahead5x6", "wide3x17")


@pytest.mark.parametrize("name", WINDOW_CASES)
def test_window_and_product_match_reference_golden(oracle, name):
    """tests/golden/lda_window.npz: the reference's own window node followed by its own matrix product, per segment."""
    z = np.load(os.path.join(GOLDEN_DIR, "lda_window.npz"))
    base, fo, m = z[name + "_base"], z[name + "_frame_offsets"], z[name + "_matrix"]
    length, right = int(z[name + "_length"]), int(z[name + "_right"])
    eye = np.eye(base.shape[1] * length, dtype=np.float32)                    # output = the window itself
    assert np.array_equal(bits(oracle.lda(base, fo, length, right, eye)), bits(z[name + "_window"]))
    assert np.array_equal(bits(oracle.lda(base, fo, length, right, m)), bits(z[name + "_out"]))


def test_window_matches_compiled_reference_node(oracle):
    """Fresh random shapes through the unmodified Signal::DelayNode (segments shorter than the window included)."""
    if not oracle.ref_window_available():
        pytest.skip("oracle/_ref/ref_window not built")
    rng = np.random.default_rng(7)
    for _ in range(40):
        T, bd, length = int(rng.integers(1, 40)), int(rng.integers(1, 9)), int(rng.integers(1, 14))
        right = int(rng.integers(0, length))
        base = rng.standard_normal((T, bd)).astype(np.float32)
        want = oracle.run_ref_window(base, length, right)
        got = oracle.lda(base, [0, T], length, right, np.eye(bd * length, dtype=np.float32))
        assert want.shape == got.shape and np.array_equal(bits(got), bits(want)), (T, bd, length, right)
