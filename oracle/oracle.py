"""TEST INFRASTRUCTURE -- NOT PRODUCT CODE.

ctypes front end of oracle/gmm_oracle.c (the CPU restatement of the reference's scorers) and a runner for
oracle/_ref/ref_scorer (the unmodified reference classes compiled by oracle/ref_build/Makefile).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgmm_oracle.so")
REF_BIN = os.path.join(HERE, "_ref", "ref_scorer")
REFERENCE_ROOT = "/root/reference"


def build(force: bool = False) -> str:
    """gcc the C restatement (no -ffast-math: operation order is explicit in the source)."""
    src = os.path.join(HERE, "gmm_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-msse3", "-fopenmp", "-fPIC", "-shared", "-funsigned-char",
                               "-o", LIB_PATH, src, "-lm"])
    return LIB_PATH


def build_ref(force: bool = False) -> str | None:
    """Compile the reference scorer classes from /root/reference (build container only)."""
    if not os.path.isdir(REFERENCE_ROOT):
        return REF_BIN if os.path.exists(REF_BIN) else None
    if force or not all(os.path.exists(os.path.join(HERE, "_ref", b)) for b in ("ref_scorer", "ref_matmul", "ref_estimator", "ref_aligner", "ref_window")):
        subprocess.check_call(["make", "-s", "-j8", "-C", os.path.join(HERE, "ref_build")])
    return REF_BIN


class _Model(C.Structure):
    _fields_ = [("D", C.c_uint32), ("M", C.c_uint32), ("nDens", C.c_uint32), ("nMeans", C.c_uint32),
                ("nCov", C.c_uint32), ("mixOffsets", C.c_void_p), ("mixDensity", C.c_void_p),
                ("mixLogWeight", C.c_void_p), ("densMean", C.c_void_p), ("densCov", C.c_void_p),
                ("means", C.c_void_p), ("variances", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.orc_simd_padded_dimension.restype = C.c_uint32
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


def _cmodel(ms):
    return _Model(ms.D, ms.M, ms.n_densities, ms.n_means, ms.n_cov, ms.mix_offsets.ctypes.data,
                  ms.mix_density.ctypes.data, ms.mix_logw.ctypes.data, ms.dens_mean.ctypes.data,
                  ms.dens_cov.ctypes.data, ms.means.ctypes.data, ms.variances.ctypes.data)


def float_tables(ms, *, mixture_weight_scale=1.0, gaussian_scale=1.0, exact_rsqrt=False):
    """(isr [C,D], logNorm [C], w2 [G]) as GaussDiagonalMaximumFeatureScorer::init builds them."""
    cm = _cmodel(ms)
    isr = np.empty((ms.n_cov, ms.D), np.float32)
    logn = np.empty(ms.n_cov, np.float32)
    w2 = np.empty(ms.G, np.float32)
    factor = np.float32(np.sqrt(np.float64(gaussian_scale)))  # ctor: gaussianScale_(std::sqrt(param)) f64->f32
    lib().orc_covariance_tables(C.byref(cm), C.c_float(float(factor)), int(exact_rsqrt), _p(isr), _p(logn))
    lib().orc_minus2_log_weights(C.byref(cm), C.c_float(mixture_weight_scale), _p(w2))
    return isr, logn, w2


def score_max(ms, x, tables=None):
    x = np.ascontiguousarray(x, np.float32)
    isr, logn, w2 = tables if tables is not None else float_tables(ms)
    T = x.shape[0]
    score = np.empty((T, ms.M), np.float32)
    best = np.empty((T, ms.M), np.uint32)
    cm = _cmodel(ms)
    lib().orc_score_max(C.byref(cm), _p(isr), _p(logn), _p(w2), _p(x), C.c_size_t(T), _p(score), _p(best))
    return score, best


def score_sum(ms, x, tables=None, posterior=False):
    x = np.ascontiguousarray(x, np.float32)
    isr, logn, w2 = tables if tables is not None else float_tables(ms)
    T = x.shape[0]
    score = np.empty((T, ms.M), np.float32)
    best = np.empty((T, ms.M), np.uint32)
    post = np.empty((T, ms.G), np.float32) if posterior else None
    cm = _cmodel(ms)
    lib().orc_score_sum(C.byref(cm), _p(isr), _p(logn), _p(w2), _p(x), C.c_size_t(T), _p(score), _p(best), _p(post))
    return (score, best, post) if posterior else (score, best)


def simd_prepare(ms, exact_rsqrt=False):
    """Quantized tables of SimdGaussDiagonalMaximumFeatureScorer::init."""
    cm = _cmodel(ms)
    Dpad = int(lib().orc_simd_padded_dimension(ms.D))
    isrs = np.empty((ms.n_cov, ms.D), np.float32)
    logns = np.empty(ms.n_cov, np.float32)
    s2 = C.c_float(0)
    qmean = np.empty((ms.G, Dpad), np.uint8)
    qconst = np.empty(ms.G, np.int32)
    qcov = np.empty(ms.G, np.uint32)
    lib().orc_simd_prepare(C.byref(cm), int(exact_rsqrt), _p(isrs), _p(logns), C.byref(s2), _p(qmean), _p(qconst),
                           _p(qcov))
    return dict(isrs=isrs, logns=logns, scaling_squared=np.float32(s2.value), qmean=qmean, qconst=qconst,
                qcov=qcov, Dpad=Dpad)


def batchint_prepare(ms, exact_rsqrt=False):
    """Quantized tables of BatchIntFeatureScorer::init (`batch-diagonal-maximum-int`); pooled covariance only."""
    cm = _cmodel(ms)
    Dpad = int(lib().orc_simd_padded_dimension(ms.D))
    isrs = np.empty((1, ms.D), np.float32)
    scale2 = C.c_float(0)
    qmean = np.empty((ms.G, Dpad), np.uint8)
    qconst = np.empty(ms.G, np.int32)
    if lib().orc_batchint_prepare(C.byref(cm), int(exact_rsqrt), _p(isrs), C.byref(scale2), _p(qmean), _p(qconst)) != 0:
        raise ValueError("feature scorer supports only globally pooled variance")
    return dict(isrs=isrs, scale2=np.float32(scale2.value), qmean=qmean, qconst=qconst,
                qcov=np.zeros(ms.G, np.uint32), Dpad=Dpad)


def batchint_final_score(q, scale2):
    q = np.ascontiguousarray(q, np.int32)
    out = np.empty(q.shape, np.float32)
    lib().orc_batchint_final_score(_p(q), C.c_size_t(q.size), C.c_float(float(scale2)), _p(out))
    return out


def simd_quantize_features(x, isrs):
    x = np.ascontiguousarray(x, np.float32)
    T, D = x.shape
    nCov = isrs.shape[0]
    Dpad = int(lib().orc_simd_padded_dimension(D))
    qx = np.empty((T, nCov, Dpad), np.uint8)
    lib().orc_simd_quantize_features(_p(x), C.c_size_t(T), D, nCov, _p(isrs), _p(qx))
    return qx


def simd_score(mix_offsets, tab, qx):
    T = qx.shape[0]
    M = mix_offsets.shape[0] - 1
    q = np.empty((T, M), np.int32)
    best = np.empty((T, M), np.uint32)
    lib().orc_simd_score(M, _p(mix_offsets), tab["Dpad"], qx.shape[1], _p(tab["qmean"]), _p(tab["qconst"]),
                         _p(tab["qcov"]), _p(qx), C.c_size_t(T), _p(q), _p(best))
    return q, best


def simd_final_score(q, scaling_squared):
    q = np.ascontiguousarray(q, np.int32)
    out = np.empty(q.shape, np.float32)
    lib().orc_simd_final_score(_p(q), C.c_size_t(q.size), C.c_float(float(scaling_squared)), _p(out))
    return out


# ---------------------------------------------------------------------------------------------------
# Runner for the compiled reference (oracle/_ref/ref_scorer)
# ---------------------------------------------------------------------------------------------------
def viterbi_align(frame_offsets, scores, automata):
    """orc_viterbi_align per segment.  `automata` has the fields of rasr_b200.align.AlignmentAutomata.
    -> (path_emission [frames] u32, path_state [frames] u32, path_score [segments] f32)."""
    L = lib()
    L.orc_viterbi_align.restype = C.c_float
    fo = np.ascontiguousarray(frame_offsets, np.uint32)
    scores = np.ascontiguousarray(scores, np.float32)
    a = automata
    n_seg = fo.shape[0] - 1
    em = np.empty(int(fo[-1]), np.uint32)
    st = np.empty(int(fo[-1]), np.uint32)
    sc = np.empty(n_seg, np.float32)
    ld = scores.shape[1] if scores.ndim == 2 else 1
    for i in range(n_seg):
        f0, f1 = int(fo[i]), int(fo[i + 1])
        s0, s1 = int(a.state_offsets[i]), int(a.state_offsets[i + 1])
        a0 = int(a.arc_offsets[s0])
        ao = np.ascontiguousarray(a.arc_offsets[s0:s1 + 1] - np.uint32(a0))
        tg = np.ascontiguousarray(a.arc_target[a0:])
        ae = np.ascontiguousarray(a.arc_emission[a0:])
        aw = np.ascontiguousarray(a.arc_weight[a0:])
        fw = np.ascontiguousarray(a.final_weight[s0:s1])
        rows = scores[f0:f1]
        e_i = np.empty(f1 - f0, np.uint32)
        s_i = np.empty(f1 - f0, np.uint32)
        sc[i] = L.orc_viterbi_align(C.c_uint32(f1 - f0), _p(rows) if f1 > f0 else C.c_void_p(0), C.c_size_t(ld),
                                    C.c_uint32(s1 - s0), _p(ao), _p(tg), _p(ae), _p(aw),
                                    C.c_uint32(int(a.initial_state[i])), _p(fw), _p(e_i), _p(s_i))
        em[f0:f1] = e_i
        st[f0:f1] = s_i
    return em, st, sc


def accumulate(ms, x, mixture, weight=None, *, sum_scorer=False, viterbi=True, weight_threshold=0.0, tables=None):
    """orc_accumulate: the reference's E-step in frame order.
    -> dict(mean_sum [nMeans, D], mean_weight, cov_sum [nCov, D], cov_weight, mix_weight [G], best [T])."""
    isr, logn, w2 = tables if tables is not None else float_tables(ms)
    x = np.ascontiguousarray(x, np.float32)
    mixture = np.ascontiguousarray(mixture, np.uint32)
    weight = None if weight is None else np.ascontiguousarray(weight, np.float64)
    T, D = x.shape
    n_means, n_cov, G = ms.n_means, ms.n_cov, ms.G
    out = dict(mean_sum=np.zeros((n_means, D)), mean_weight=np.zeros(n_means), cov_sum=np.zeros((n_cov, D)),
               cov_weight=np.zeros(n_cov), mix_weight=np.zeros(G), best=np.empty(T, np.uint32))
    cm = _cmodel(ms)
    lib().orc_accumulate(C.byref(cm), _p(isr), _p(logn), _p(w2), _p(x), C.c_size_t(T), _p(mixture), _p(weight),
                         C.c_int(int(sum_scorer)), C.c_int(int(viterbi)), C.c_double(weight_threshold),
                         _p(out["mean_sum"]), _p(out["mean_weight"]), _p(out["cov_sum"]), _p(out["cov_weight"]),
                         _p(out["mix_weight"]), _p(out["best"]))
    return out


def lda(base, frame_offsets, length, right, matrix):
    """orc_lda: window (length frames, `right` after the current one, clamped to the segment) + matrix product."""
    base = np.ascontiguousarray(base, np.float32)
    matrix = np.ascontiguousarray(matrix, np.float32)
    fo = np.ascontiguousarray(frame_offsets, np.uint32)
    T, base_dim = base.shape
    D = matrix.shape[0]
    assert matrix.shape[1] == base_dim * length
    out = np.empty((T, D), np.float32)
    lib().orc_lda(_p(base), C.c_uint32(fo.shape[0] - 1), _p(fo), C.c_uint32(base_dim), C.c_uint32(length),
                  C.c_uint32(right), _p(matrix), C.c_uint32(D), _p(out))
    return out


def matrix_vector(matrix, vectors):
    matrix = np.ascontiguousarray(matrix, np.float32)
    vectors = np.ascontiguousarray(vectors, np.float32)
    out = np.empty((vectors.shape[0], matrix.shape[0]), np.float32)
    for i in range(vectors.shape[0]):
        lib().orc_matrix_vector(_p(matrix), C.c_uint32(matrix.shape[0]), C.c_uint32(matrix.shape[1]),
                                _p(vectors[i]), _p(out[i]))
    return out


REF_MATMUL = os.path.join(HERE, "_ref", "ref_matmul")


def run_ref_matmul(matrix, vectors):
    """The reference's Math::Matrix<f32> * Math::Vector<f32> (oracle/_ref/ref_matmul)."""
    matrix = np.ascontiguousarray(matrix, np.float32)
    vectors = np.ascontiguousarray(vectors, np.float32)
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "in"), "wb") as f:
            f.write(np.array([matrix.shape[0], matrix.shape[1], vectors.shape[0]], np.uint32).tobytes())
            f.write(matrix.tobytes())
            f.write(vectors.tobytes())
        subprocess.check_call([REF_MATMUL, os.path.join(d, "in"), os.path.join(d, "out")])
        return np.fromfile(os.path.join(d, "out"), np.float32).reshape(vectors.shape[0], matrix.shape[0])


def ref_available() -> bool:
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def run_ref(ms, x, scorer="max", *, threads=0, repeat=1, dump=True, workdir=None):
    """Run the unmodified reference scorer `scorer` (max|sum|simd|batchint|batchfloat) on (ms, x).

    Returns a dict of numpy arrays (score, best, and for simd q/qmean/qconst/qcov/isrs/qx/meta) plus
    "timing" when threads > 0."""
    if not ref_available():
        raise RuntimeError("oracle/_ref/ref_scorer is not built (needs /root/reference: oracle.build_ref())")
    x = np.ascontiguousarray(x, np.float32)
    T, M = x.shape[0], ms.M
    with tempfile.TemporaryDirectory(dir=workdir) as tmp:
        mp, fp, op = os.path.join(tmp, "model.gms"), os.path.join(tmp, "feat.f32"), os.path.join(tmp, "out")
        ms.save(mp)
        x.tofile(fp)
        cmd = [REF_BIN, f"model={mp}", f"features={fp}", f"scorer={scorer}", f"out={op}",
               f"dump={1 if dump else 0}", f"threads={threads}", f"repeat={repeat}"]
        r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"ref_scorer failed ({r.returncode}):\n{r.stdout[-2000:]}")
        out = {"log": r.stdout}
        if dump:
            out["score"] = np.fromfile(op + ".score.f32", np.float32).reshape(T, M)
            if os.path.exists(op + ".best.u32"):
                out["best"] = np.fromfile(op + ".best.u32", np.uint32).reshape(T, M)
            if scorer == "simd":
                meta = {}
                with open(op + ".meta.txt") as f:
                    for line in f:
                        k, v = line.split()
                        meta[k] = float(v) if k == "scalingSquared" else int(v)
                Dpad, G, Cn = meta["Dpad"], meta["G"], meta["C"]
                out["meta"] = meta
                out["scaling_squared"] = np.float32(meta["scalingSquared"])
                out["q"] = np.fromfile(op + ".q.i32", np.int32).reshape(T, M)
                out["qmean"] = np.fromfile(op + ".qmean.u8", np.uint8).reshape(G, Dpad)
                out["qconst"] = np.fromfile(op + ".qconst.i32", np.int32)
                out["qcov"] = np.fromfile(op + ".qcov.u32", np.uint32)
                out["isrs"] = np.fromfile(op + ".isrs.f32", np.float32).reshape(Cn, ms.D)
                out["qx"] = np.fromfile(op + ".qx.u8", np.uint8).reshape(T, Cn, Dpad)
        if threads > 0:
            with open(op + ".timing.json") as f:
                out["timing"] = json.load(f)
        return out


REF_WINDOW_BIN = os.path.join(HERE, "_ref", "ref_window")


def ref_window_available() -> bool:
    return os.path.exists(REF_WINDOW_BIN) and os.access(REF_WINDOW_BIN, os.X_OK)


def run_ref_window(base, length, right, *, margin_policy="copy", margin_condition="present-not-empty", workdir=None):
    """The reference's sliding-window node (Signal::DelayNode through signal-vector-f32-sequence-concatenation, wired as
    share/lda.flow wires it; oracle/ref_build/ref_window.cc) over ONE segment `base` [frames x baseDim].
    -> [outputs x baseDim*length] f32."""
    if not ref_window_available():
        raise RuntimeError("oracle/_ref/ref_window is not built (needs /root/reference: oracle.build_ref())")
    base = np.ascontiguousarray(base, np.float32)
    with tempfile.TemporaryDirectory(dir=workdir) as tmp:
        ip, op = os.path.join(tmp, "in.bin"), os.path.join(tmp, "out.bin")
        with open(ip, "wb") as f:
            np.asarray([base.shape[0], base.shape[1], length, right], np.uint32).tofile(f)
            base.tofile(f)
        r = subprocess.run([REF_WINDOW_BIN, f"in={ip}", f"out={op}", f"margin-policy={margin_policy}",
                            f"margin-condition={margin_condition}"], cwd=tmp, stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"ref_window failed ({r.returncode}):\n{r.stdout[-2000:]}")
        n, width = np.fromfile(op, np.uint32, 2)
        return np.fromfile(op, np.float32, offset=8).reshape(int(n), int(width)) if n else np.zeros((0, base.shape[1] * length), np.float32)


REF_EST_BIN = os.path.join(HERE, "_ref", "ref_estimator")


def ref_estimator_available() -> bool:
    return os.path.exists(REF_EST_BIN) and os.access(REF_EST_BIN, os.X_OK)


def run_ref_estimator(ms, x=None, mixture=None, weight=None, *, mode="viterbi", scorer="max", min_variance=None,
                      threshold=None, workdir=None):
    """Run the unmodified reference text writer / reader and estimator stack (oracle/ref_build/ref_estimator.cc).
    -> dict(pms, reread_pms [, best, mixset, mixset2, est_pms]): the files as bytes (best: u32 array)."""
    if not ref_estimator_available():
        raise RuntimeError("oracle/_ref/ref_estimator is not built (needs /root/reference: oracle.build_ref())")
    with tempfile.TemporaryDirectory(dir=workdir) as tmp:
        mp, op = os.path.join(tmp, "model.gms"), os.path.join(tmp, "out")
        ms.save(mp)
        cmd = [REF_EST_BIN, f"model={mp}", f"out={op}"]
        if x is not None:
            fp, ap = os.path.join(tmp, "feat.f32"), os.path.join(tmp, "align.u32")
            np.ascontiguousarray(x, np.float32).tofile(fp)
            np.ascontiguousarray(mixture, np.uint32).tofile(ap)
            cmd += [f"features={fp}", f"align={ap}", f"mode={mode}", f"scorer={scorer}"]
            if weight is not None:
                wp = os.path.join(tmp, "weight.f64")
                np.ascontiguousarray(weight, np.float64).tofile(wp)
                cmd.append(f"weights={wp}")
            if min_variance is not None:
                cmd.append(f"min-variance={min_variance!r}")
            if threshold is not None:
                cmd.append(f"threshold={threshold!r}")
        r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"ref_estimator failed ({r.returncode}):\n{r.stdout[-2000:]}")
        out = {"log": r.stdout}
        for key, ext in (("pms", ".pms"), ("reread_pms", ".reread.pms"), ("mixset", ".mixset"), ("mixset2", ".mixset2"),
                         ("est_pms", ".est.pms")):
            if os.path.exists(op + ext):
                with open(op + ext, "rb") as f:
                    out[key] = f.read()
        if os.path.exists(op + ".best.u32"):
            out["best"] = np.fromfile(op + ".best.u32", np.uint32)
        return out


REF_ALIGNER_BIN = os.path.join(HERE, "_ref", "ref_aligner")


def ref_aligner_available() -> bool:
    return os.path.exists(REF_ALIGNER_BIN) and os.access(REF_ALIGNER_BIN, os.X_OK)


def run_ref_aligner(scores, automata, i=0, *, mode="viterbi", emission_of_label=None, workdir=None):
    """Run the unmodified reference Search::Aligner (oracle/ref_build/ref_aligner.cc) on segment i's automaton over the
    score rows `scores` [frames x emissions], unpruned.  With emission_of_label the arcs' "emission" field is a label
    the acoustic model maps to an emission (state tying); the items then name labels.
    -> dict(time u32 [n], emission u32 [n], weight f64 [n], score f32 (Aligner::alignmentScore()),
            total f32 (baum-welch: totalInv of Fsa::posterior64 = MINUS the total, Aligner.cc:757-772))."""
    if not ref_aligner_available():
        raise RuntimeError("oracle/_ref/ref_aligner is not built (needs /root/reference: oracle.build_ref())")
    a = automata
    scores = np.ascontiguousarray(scores, np.float32)
    s0, s1 = int(a.state_offsets[i]), int(a.state_offsets[i + 1])
    a0, a1 = int(a.arc_offsets[s0]), int(a.arc_offsets[s1])
    with tempfile.TemporaryDirectory(dir=workdir) as tmp:
        ip, op = os.path.join(tmp, "in.bin"), os.path.join(tmp, "out")
        with open(ip, "wb") as f:
            np.asarray([scores.shape[0], scores.shape[1], s1 - s0, a1 - a0, int(a.initial_state[i])], np.uint32).tofile(f)
            np.ascontiguousarray(a.arc_offsets[s0:s1 + 1] - np.uint32(a0), np.uint32).tofile(f)
            np.ascontiguousarray(a.arc_target[a0:a1], np.uint32).tofile(f)
            np.ascontiguousarray(a.arc_emission[a0:a1], np.uint32).tofile(f)
            np.ascontiguousarray(a.arc_weight[a0:a1], np.float32).tofile(f)
            np.ascontiguousarray(a.final_weight[s0:s1], np.float32).tofile(f)
            scores.tofile(f)
        cmd = [REF_ALIGNER_BIN, f"in={ip}", f"out={op}", f"mode={mode}", "--*.aligner.min-acoustic-pruning=1e30",
               "--*.aligner.max-acoustic-pruning=1e30"]
        if emission_of_label is not None:   # arcs carry labels; the acoustic model ties label -> emission
            lp = os.path.join(tmp, "labels.bin")
            with open(lp, "wb") as f:
                np.asarray([len(emission_of_label)], np.uint32).tofile(f)
                np.ascontiguousarray(emission_of_label, np.uint32).tofile(f)
            cmd.append(f"labels={lp}")
        r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"ref_aligner failed ({r.returncode}):\n{r.stdout[-2000:]}")
        with open(op + ".items.bin", "rb") as f:
            raw = f.read()
        n = int(np.frombuffer(raw[:4], np.uint32)[0])
        items = np.frombuffer(raw[4:], np.dtype([("t", "<u4"), ("e", "<u4"), ("w", "<f8")]), count=n)
        sc = np.fromfile(op + ".score.f32", np.float32)
    return {"time": items["t"].copy(), "emission": items["e"].copy(), "weight": items["w"].copy(), "score": sc[0],
            "total": sc[1], "log": r.stdout}


def emission_lists(automata, i):
    """Distinct emissions of segment i's automaton in order of first use by an arc (column order of the compact
    score / posterior blocks, include/gmm_b200.h)."""
    a = automata
    s0, s1 = int(a.state_offsets[i]), int(a.state_offsets[i + 1])
    a0, a1 = int(a.arc_offsets[s0]), int(a.arc_offsets[s1])
    seen, out = set(), []
    for e in a.arc_emission[a0:a1].tolist():
        if e not in seen:
            seen.add(e)
            out.append(e)
    return np.asarray(out, np.uint32)


def baum_welch_align(frame_offsets, scores, automata):
    """orc_baum_welch_align per segment -> (weights: list of [frames_i x len(list_i)] f32, lists: list of u32 arrays,
    total [segments] f64)."""
    L = lib()
    L.orc_baum_welch_align.restype = C.c_double
    fo = np.ascontiguousarray(frame_offsets, np.uint32)
    scores = np.ascontiguousarray(scores, np.float32)
    a = automata
    n_seg = fo.shape[0] - 1
    ld = scores.shape[1] if scores.ndim == 2 else 1
    weights, lists, total = [], [], np.empty(n_seg, np.float64)
    for i in range(n_seg):
        f0, f1 = int(fo[i]), int(fo[i + 1])
        s0, s1 = int(a.state_offsets[i]), int(a.state_offsets[i + 1])
        a0 = int(a.arc_offsets[s0])
        ao = np.ascontiguousarray(a.arc_offsets[s0:s1 + 1] - np.uint32(a0))
        tg = np.ascontiguousarray(a.arc_target[a0:])
        ae = np.ascontiguousarray(a.arc_emission[a0:])
        aw = np.ascontiguousarray(a.arc_weight[a0:])
        fw = np.ascontiguousarray(a.final_weight[s0:s1])
        lst = emission_lists(a, i)
        rows = scores[f0:f1]
        w = np.zeros((f1 - f0, max(1, lst.shape[0])), np.float32)
        total[i] = L.orc_baum_welch_align(C.c_uint32(f1 - f0), _p(rows) if f1 > f0 else C.c_void_p(0), C.c_size_t(ld),
                                          C.c_uint32(s1 - s0), _p(ao), _p(tg), _p(ae), _p(aw),
                                          C.c_uint32(int(a.initial_state[i])), _p(fw), C.c_uint32(lst.shape[0]),
                                          _p(lst) if lst.shape[0] else C.c_void_p(0), _p(w), C.c_size_t(w.shape[1]))
        weights.append(w[:, :lst.shape[0]])
        lists.append(lst)
    return weights, lists, total
