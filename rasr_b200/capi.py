"""ctypes binding of the C-ABI in include/gmm_b200.h (libgmm_b200.so, hand-written CUDA for sm_100a).

This is the binding a maintainer would write for any host language; it is what the tests, bench.py and the
Python mirror of the reference's scorer interface (rasr_b200/feature_scorer.py) call.  There is NO fallback:
if the shared library is missing, or no sm_100 device is present, loading / ``Engine()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GMM_B200_LIB") or os.path.join(HERE, "libgmm_b200.so")  # (override: kernel experiments)

MODE_MAX, MODE_SUM = 0, 1
KERNEL_AUTO, KERNEL_DIRECT, KERNEL_TENSOR = 0, 1, 2
QUANT_RULE_SIMD, QUANT_RULE_BATCH_INT = 0, 1  # gmm_quant_final_score_rule
QUANT_STATIC_SCRATCH_BYTES = 4 << 20  # GMM_QUANT_STATIC_SCRATCH_BYTES

# every symbol include/gmm_b200.h declares: (name, restype, argtypes)
_vp, _u32, _sz, _i = C.c_void_p, C.c_uint32, C.c_size_t, C.c_int
SYMBOLS = [
    ("gmm_create", _i, [C.POINTER(_vp), _i]),
    ("gmm_destroy", None, [_vp]),
    ("gmm_last_error", C.c_char_p, [_vp]),
    ("gmm_set_stream", _i, [_vp, _vp]),
    ("gmm_upload_float", _i, [_vp, _u32, _u32, _u32, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i]),
    ("gmm_score_host", _i, [_vp, _u32, _vp, _vp, _vp, _sz]),
    ("gmm_score_async", _i, [_vp, _u32, _vp, _vp, _vp, _sz, C.POINTER(C.c_uint64)]),
    ("gmm_wait", _i, [_vp, C.c_uint64]),
    ("gmm_score_device", _i, [_vp, _u32, _vp, _vp, _vp, _sz]),
    ("gmm_posteriors_device", _i, [_vp, _u32, _vp, _vp]),
    ("gmm_mixture_posteriors_host", _i, [_vp, _u32, _vp, _vp, _vp, _sz, _vp]),
    ("gmm_upload_quant", _i, [_vp, _u32, _u32, _u32, _u32, _u32, _vp, _vp, _vp, _vp, _vp, C.c_float]),
    ("gmm_quant_score_host", _i, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _sz]),
    ("gmm_quant_score_device", _i, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _sz]),
    ("gmm_quant_score_device_static", _i, [_vp, _u32, _vp, _vp, _vp, _vp, _sz]),
    ("gmm_emission_lists", _i, [_vp, _u32, _vp, _vp, _vp, _u32, _vp, _vp]),
    ("gmm_baum_welch_device", _i, [_vp, _u32, _vp, _vp, _sz, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    ("gmm_score_baum_welch_host", _i, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    ("gmm_score_baum_welch_accumulate_host", _i, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    ("gmm_score_device_gather", _i, [_vp, _u32, _vp, _u32, _u32, _vp, _sz, _u32, _sz, _u32]),
    ("gmm_score_device_gather_begin", _i, [_vp, _u32, _vp, _u32, _u32, _vp, _sz, _u32, _sz, _u32]),
    ("gmm_gather_finish", _i, [_vp, _vp, _sz]),
    ("gmm_gather_status", _i, [_vp, _vp]),
    ("gmm_gather_reset", _i, [_vp]),
    ("gmm_peer_alloc", _i, [_vp, _sz, _vp, _vp]),
    ("gmm_peer_open", _i, [_vp, _vp, _vp]),
    ("gmm_peer_close", _i, [_vp, _vp]),
    ("gmm_peer_free", _i, [_vp, _vp]),
    ("gmm_score_subset_device", _i, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _sz]),
    ("gmm_score_subset_ld", _u32, [_vp, _u32]),
    ("gmm_align_device", _i, [_vp, _u32, _vp, _vp, _sz] + [_vp] * 10),
    ("gmm_align_ordered_count", C.c_uint64, [_vp]),
    ("gmm_set_consumer_score_scale", _i, [_vp, C.c_float]),
    ("gmm_score_align_host", _i, [_vp, _u32, _vp, _vp] + [_vp] * 10),
    ("gmm_score_align_accumulate_host", _i, [_vp, _u32, _vp, _vp] + [_vp] * 10),
    ("gmm_score_align_labels_host", _i, [_vp, _u32, _vp, _vp] + [_vp] * 12),
    ("gmm_score_baum_welch_labels_host", _i, [_vp, _u32, _vp, _vp] + [_vp] * 10 + [_sz, _vp]),
    ("gmm_accumulator_create", _i, [_vp, _u32, _vp, _i, C.c_double]),
    ("gmm_accumulator_reset", _i, [_vp]),
    ("gmm_accumulate_device", _i, [_vp, _u32, _vp, _vp, _vp, _vp]),
    ("gmm_accumulate_host", _i, [_vp, _u32, _vp, _vp, _vp, _vp]),
    ("gmm_accumulator_size", _sz, [_vp]),
    ("gmm_accumulator_export_device", _i, [_vp, _vp]),
    ("gmm_accumulator_import_device", _i, [_vp, _vp]),
    ("gmm_accumulator_read", _i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    ("gmm_set_lda", _i, [_vp, _u32, _u32, _u32, _u32, _vp]),
    ("gmm_lda_device", _i, [_vp, _u32, _vp, _vp, _vp]),
    ("gmm_host_alloc", _vp, [_sz]),
    ("gmm_host_free", None, [_vp]),
    ("gmm_sync", _i, [_vp]),
    ("gmm_launch_count", C.c_uint64, [_vp]),
    ("gmm_kernel_timing", _i, [_vp, _i]),
    ("gmm_kernel_time", _i, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]),
    ("gmm_float_kernel", _i, [_vp]),
    ("gmm_float_split", _i, [_vp]),
    ("gmm_float_accuracy_ok", _i, [_vp]),
    ("gmm_quant_final_score_rule", _i, [_vp, _i]),
    ("gmm_quant_kernel", _i, [_vp]),
    ("gmm_version", C.c_char_p, []),
]

_lib = None


class GmmError(RuntimeError):
    pass


def load():
    """dlopen libgmm_b200.so and bind every declared symbol; raises if the library was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GmmError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(make -C rasr_b200/csrc).  There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if isinstance(a, int):
        return a
    return a.data_ptr()  # torch tensor


def _np(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


class Engine:
    """One gmm_handle: one GPU, one CUDA stream, at most one float and one quantized model."""

    def __init__(self, device: int = 0):
        self._lib = load()
        h = _vp()
        rc = self._lib.gmm_create(C.byref(h), int(device))
        if rc != 0:
            raise GmmError(f"gmm_create failed ({rc}): {self._lib.gmm_last_error(None).decode()}")
        self._h = h
        self.M = self.G = self.D = 0
        self.qM = 0

    def close(self):
        if getattr(self, "_h", None):
            self._lib.gmm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise GmmError(f"{what} failed ({rc}): {self._lib.gmm_last_error(self._h).decode()}")

    # -- plumbing ---------------------------------------------------------------------------------
    def set_stream(self, cuda_stream_ptr: int | None):
        self._check(self._lib.gmm_set_stream(self._h, cuda_stream_ptr), "gmm_set_stream")

    def sync(self):
        self._check(self._lib.gmm_sync(self._h), "gmm_sync")

    @property
    def launch_count(self) -> int:
        return int(self._lib.gmm_launch_count(self._h))

    def kernel_timing(self, on: bool):
        self._check(self._lib.gmm_kernel_timing(self._h, 1 if on else 0), "gmm_kernel_timing")

    def kernel_time(self):
        """(summed device ms, launches) of the scoring kernels bracketed since the last call."""
        ms, n = C.c_double(0.0), C.c_uint64(0)
        self._check(self._lib.gmm_kernel_time(self._h, C.byref(ms), C.byref(n)), "gmm_kernel_time")
        return ms.value, int(n.value)

    @property
    def float_kernel(self) -> int:
        return int(self._lib.gmm_float_kernel(self._h))

    @property
    def float_split(self) -> int:
        """3: three fp16 products (default); 2: fp16 + e5m2 corrections (GMM_TC_SPLIT=fp8); 0: not on the tensor kernel."""
        return int(self._lib.gmm_float_split(self._h))

    @property
    def float_accuracy_ok(self) -> bool:
        """False only when the tensor kernel was forced for a model the automatic choice sends to the direct kernel."""
        return bool(self._lib.gmm_float_accuracy_ok(self._h))

    @property
    def quant_kernel(self) -> int:
        return int(self._lib.gmm_quant_kernel(self._h))

    # -- float engine -----------------------------------------------------------------------------
    def upload_float(self, offsets, mean, isr, cov_idx, w2, log_norm, mode=MODE_MAX, kernel=KERNEL_AUTO):
        offsets = _np(offsets, np.uint32)
        mean = _np(mean, np.float32)
        isr = _np(isr, np.float32)
        cov_idx = _np(cov_idx, np.uint32)
        w2 = _np(w2, np.float32)
        log_norm = _np(log_norm, np.float32)
        M, G = offsets.shape[0] - 1, int(offsets[-1])
        Cn, D = isr.shape
        if mean.shape != (G, D) or cov_idx.shape != (G,) or w2.shape != (G,) or log_norm.shape != (Cn,):
            raise ValueError("upload_float: table shapes are inconsistent")
        self._check(self._lib.gmm_upload_float(self._h, M, G, D, Cn, _ptr(offsets), _ptr(mean), _ptr(isr),
                                               _ptr(cov_idx), _ptr(w2), _ptr(log_norm), mode, kernel),
                    "gmm_upload_float")
        self.M, self.G, self.D = M, G, D

    def score_host(self, x, want_best=True, ld=None, out=None):
        """x [T, D] f32 host array -> (scores [T, M] f32, best [T, M] u32 | None), host arrays."""
        x = _np(x, np.float32)
        if x.ndim != 2 or x.shape[1] != self._in_dim(self.D):
            raise ValueError(f"features must be [T, {self._in_dim(self.D)}]")
        T = x.shape[0]
        ld = self.M if ld is None else int(ld)
        scores = out if out is not None else np.empty((T, ld), np.float32)
        best = np.empty((T, ld), np.uint32) if want_best else None
        self._check(self._lib.gmm_score_host(self._h, T, _ptr(x), _ptr(scores), _ptr(best), ld), "gmm_score_host")
        return scores, best

    def score_async(self, x, scores, best=None, ld=None) -> int:
        """gmm_score_async: x [T, D], scores [T, ld] (and best) are HOST arrays the caller keeps alive (page-locked for
        real overlap) until wait(token); returns the token."""
        T = x.shape[0]
        ld = self.M if ld is None else int(ld)
        tok = C.c_uint64(0)
        self._check(self._lib.gmm_score_async(self._h, int(T), _ptr(x), _ptr(scores), _ptr(best), ld, C.byref(tok)),
                    "gmm_score_async")
        return int(tok.value)

    def wait(self, token: int):
        self._check(self._lib.gmm_wait(self._h, C.c_uint64(int(token))), "gmm_wait")

    def score_device(self, T, d_x, d_scores, d_best=None, ld=None):
        """Device pointers (ints or torch tensors); asynchronous on the handle's stream."""
        ld = self.M if ld is None else int(ld)
        self._check(self._lib.gmm_score_device(self._h, int(T), _ptr(d_x), _ptr(d_scores), _ptr(d_best), ld),
                    "gmm_score_device")

    def posteriors_device(self, T, d_x, d_post):
        self._check(self._lib.gmm_posteriors_device(self._h, int(T), _ptr(d_x), _ptr(d_post)),
                    "gmm_posteriors_device")

    def mixture_posteriors_host(self, x, mixture, max_k: int):
        """Sum scorer: density posteriors of mixture[t] for frame t -> (post [T, max_k] f32, score [T] f32)."""
        x = _np(x, np.float32)
        mixture = _np(mixture, np.uint32)
        T = x.shape[0]
        post = np.zeros((T, int(max_k)), np.float32)
        score = np.zeros(T, np.float32)
        self._check(self._lib.gmm_mixture_posteriors_host(self._h, T, _ptr(x), _ptr(mixture), _ptr(post), int(max_k),
                                                          _ptr(score)), "gmm_mixture_posteriors_host")
        return post, score

    def subset_ld(self, max_list_length: int) -> int:
        return int(self._lib.gmm_score_subset_ld(self._h, int(max_list_length)))

    def score_subset_device(self, frame_offsets, list_offsets, mixtures, d_x, d_scores, ld):
        """Per segment, only the listed mixtures: column j of the segment's rows = mixture list[j]."""
        fo, lo, mx = _np(frame_offsets, np.uint32), _np(list_offsets, np.uint32), _np(mixtures, np.uint32)
        if lo.shape != fo.shape or int(lo[-1]) != mx.shape[0]:
            raise ValueError("list offsets must have one entry per segment boundary and end at len(mixtures)")
        self._check(self._lib.gmm_score_subset_device(self._h, fo.shape[0] - 1, _ptr(fo), _ptr(lo), _ptr(mx), _ptr(d_x),
                                                      _ptr(d_scores), int(ld)), "gmm_score_subset_device")

    # -- alignment consumer (scores never leave the device) ---------------------------------------
    def _align_args(self, frame_offsets, automata):
        fo = _np(frame_offsets, np.uint32)
        a = automata
        n_seg = fo.shape[0] - 1
        if a.state_offsets.shape[0] != n_seg + 1 or a.initial_state.shape[0] != n_seg:
            raise ValueError("align: automata and frame offsets describe different numbers of segments")
        n_frames = int(fo[-1]) if n_seg >= 0 and fo.shape[0] else 0
        return fo, n_seg, n_frames

    def align_device(self, frame_offsets, d_scores, automata, ld=None, want_states=True):
        """Best Viterbi path of every segment through its automaton over the DEVICE score block.
        -> (path_emission [frames] u32, path_state [frames] u32 | None, path_score [segments] f32), host arrays."""
        fo, n_seg, n_frames = self._align_args(frame_offsets, automata)
        a = automata
        ld = self.M if ld is None else int(ld)
        em = np.empty(n_frames, np.uint32)
        st = np.empty(n_frames, np.uint32) if want_states else None
        sc = np.empty(max(n_seg, 0), np.float32)
        self._check(self._lib.gmm_align_device(self._h, n_seg, _ptr(fo), _ptr(d_scores), ld, _ptr(a.state_offsets),
                                               _ptr(a.arc_offsets), _ptr(a.arc_target), _ptr(a.arc_emission),
                                               _ptr(a.arc_weight), _ptr(a.initial_state), _ptr(a.final_weight),
                                               _ptr(em), _ptr(st), _ptr(sc)), "gmm_align_device")
        return em, st, sc

    def score_align_host(self, frame_offsets, x, automata, want_states=True, accumulate=False, want_paths=True):
        """Host features [frames, D] -> float scorer -> alignment; only the paths come back.
        accumulate=True: the aligned frames also go into the estimators (create_accumulators), on the device;
        want_paths=False then returns nothing at all."""
        fo, n_seg, n_frames = self._align_args(frame_offsets, automata)
        a = automata
        x = _np(x, np.float32)
        if x.ndim != 2 or x.shape != (n_frames, self._in_dim(self.D)):
            raise ValueError(f"features must be [{n_frames}, {self._in_dim(self.D)}]")
        if not want_paths and not accumulate:
            raise ValueError("want_paths=False only makes sense with accumulate=True")
        em = np.empty(n_frames, np.uint32) if want_paths else None
        st = np.empty(n_frames, np.uint32) if want_states and want_paths else None
        sc = np.empty(max(n_seg, 0), np.float32) if want_paths else None
        fn = self._lib.gmm_score_align_accumulate_host if accumulate else self._lib.gmm_score_align_host
        self._check(fn(self._h, n_seg, _ptr(fo), _ptr(x), _ptr(a.state_offsets),
                                                   _ptr(a.arc_offsets), _ptr(a.arc_target), _ptr(a.arc_emission),
                                                   _ptr(a.arc_weight), _ptr(a.initial_state), _ptr(a.final_weight),
                                                   _ptr(em), _ptr(st), _ptr(sc)), "gmm_score_align_host")
        return em, st, sc

    # -- mixture-sharded variant: scoring with its exchange step -------------------------------------
    def peer_alloc(self, nbytes: int):
        """Zeroed device memory other processes of this box can map -> (device pointer, 64-byte IPC handle)."""
        ptr = C.c_void_p(0)
        handle = C.create_string_buffer(64)
        self._check(self._lib.gmm_peer_alloc(self._h, int(nbytes), C.addressof(ptr), C.addressof(handle)),
                    "gmm_peer_alloc")
        return int(ptr.value), bytes(handle.raw)

    def peer_open(self, handle: bytes) -> int:
        ptr = C.c_void_p(0)
        buf = C.create_string_buffer(handle, 64)
        self._check(self._lib.gmm_peer_open(self._h, C.addressof(buf), C.addressof(ptr)), "gmm_peer_open")
        return int(ptr.value)

    def peer_close(self, ptr: int):
        self._check(self._lib.gmm_peer_close(self._h, int(ptr)), "gmm_peer_close")

    def peer_free(self, ptr: int):
        self._check(self._lib.gmm_peer_free(self._h, int(ptr)), "gmm_peer_free")

    def score_device_gather(self, T, d_x, rank, blocks, ld, col_offset, flags_offset, chunks=4):
        """Score the shard into the own block and push its columns into the peers' blocks (stream-ordered)."""
        arr = (C.c_void_p * len(blocks))(*[int(p) for p in blocks])
        self._check(self._lib.gmm_score_device_gather(self._h, int(T), _ptr(d_x), int(rank), len(blocks),
                                                      C.addressof(arr), int(ld), int(col_offset), int(flags_offset),
                                                      int(chunks)), "gmm_score_device_gather")

    def score_device_gather_begin(self, T, d_x, rank, blocks, ld, col_offset, flags_offset, chunks=4):
        """score_device_gather without the join: the last range's push is still running when this returns."""
        arr = (C.c_void_p * len(blocks))(*[int(p) for p in blocks])
        self._check(self._lib.gmm_score_device_gather_begin(self._h, int(T), _ptr(d_x), int(rank), len(blocks),
                                                            C.addressof(arr), int(ld), int(col_offset), int(flags_offset),
                                                            int(chunks)), "gmm_score_device_gather_begin")

    def gather_finish(self, own_block: int, flags_offset: int):
        """The join of a started call: work queued on the engine's stream afterwards sees every rank's columns."""
        self._check(self._lib.gmm_gather_finish(self._h, int(own_block), int(flags_offset)), "gmm_gather_finish")

    def gather_status(self) -> int:
        """Number of the first mixture-sharded call on this handle whose exchange timed out (0: none).  Meaningful
        after sync()."""
        v = C.c_uint32(0)
        self._check(self._lib.gmm_gather_status(self._h, C.addressof(v)), "gmm_gather_status")
        return int(v.value)

    def gather_reset(self):
        """Collective by contract: every rank, between two host barriers."""
        self._check(self._lib.gmm_gather_reset(self._h), "gmm_gather_reset")

    # -- Baum-Welch alignment ----------------------------------------------------------------------
    def emission_lists(self, automata, n_emissions):
        """-> (list_offsets [segments + 1], lists): the distinct emissions of every automaton in order of first use,
        the column order of compact score blocks and of the Baum-Welch posterior rows."""
        a = automata
        n_seg = a.segments
        lo = np.empty(n_seg + 1, np.uint32)
        lst = np.empty(max(1, int(a.arc_emission.shape[0])), np.uint32)
        self._check(self._lib.gmm_emission_lists(self._h, n_seg, _ptr(a.state_offsets), _ptr(a.arc_offsets),
                                                 _ptr(a.arc_emission), int(n_emissions), _ptr(lo), _ptr(lst)),
                    "gmm_emission_lists")
        return lo, lst[:int(lo[-1])]

    def baum_welch_device(self, frame_offsets, d_scores, automata, d_weights, ld=None, ld_w=None):
        """Scores [frames, ld] and weights [frames, ld_w] on the device -> total score per segment (f64)."""
        fo, n_seg, n_frames = self._align_args(frame_offsets, automata)
        a = automata
        total = np.empty(max(n_seg, 0), np.float64)
        self._check(self._lib.gmm_baum_welch_device(
            self._h, n_seg, _ptr(fo), _ptr(d_scores), int(ld if ld is not None else d_scores.shape[1]),
            _ptr(a.state_offsets), _ptr(a.arc_offsets), _ptr(a.arc_target), _ptr(a.arc_emission), _ptr(a.arc_weight),
            _ptr(a.initial_state), _ptr(a.final_weight), _ptr(d_weights),
            int(ld_w if ld_w is not None else d_weights.shape[1]), _ptr(total)), "gmm_baum_welch_device")
        return total

    def score_baum_welch_host(self, frame_offsets, x, automata, ld_w=None, accumulate=False, want_weights=True):
        """Host features -> posterior rows [frames, ld_w] (f32, host) and total scores; accumulate=True with
        want_weights=False is the Baum-Welch training pass (nothing but the totals comes back)."""
        fo, n_seg, n_frames = self._align_args(frame_offsets, automata)
        a = automata
        x = _np(x, np.float32)
        if x.ndim != 2 or x.shape != (n_frames, self._in_dim(self.D)):
            raise ValueError(f"features must be [{n_frames}, {self._in_dim(self.D)}]")
        total = np.empty(max(n_seg, 0), np.float64)
        if accumulate:
            self._check(self._lib.gmm_score_baum_welch_accumulate_host(
                self._h, n_seg, _ptr(fo), _ptr(x), _ptr(a.state_offsets), _ptr(a.arc_offsets), _ptr(a.arc_target),
                _ptr(a.arc_emission), _ptr(a.arc_weight), _ptr(a.initial_state), _ptr(a.final_weight), _ptr(total)),
                "gmm_score_baum_welch_accumulate_host")
            return None, total
        if ld_w is None:
            lo, _ = self.emission_lists(a, self.M)
            ld_w = max(1, int(np.max(np.diff(lo)))) if n_seg else 1
        w = np.zeros((n_frames, int(ld_w)), np.float32)
        self._check(self._lib.gmm_score_baum_welch_host(
            self._h, n_seg, _ptr(fo), _ptr(x), _ptr(a.state_offsets), _ptr(a.arc_offsets), _ptr(a.arc_target),
            _ptr(a.arc_emission), _ptr(a.arc_weight), _ptr(a.initial_state), _ptr(a.final_weight), _ptr(w), int(ld_w),
            _ptr(total)), "gmm_score_baum_welch_host")
        return w, total

    def set_consumer_score_scale(self, scale: float):
        """The host-level alignment consumers read fl(scale * score) (Mm::FeatureScorerScaling)."""
        self._check(self._lib.gmm_set_consumer_score_scale(self._h, float(scale)), "gmm_set_consumer_score_scale")

    def score_align_labels_host(self, frame_offsets, x, automata, label_offsets, label_mixture, want_states=True):
        """As score_align_host for arcs that carry labels: automata.arc_emission holds segment-local label slots,
        label_mixture[label_offsets[s] + slot] the mixture scoring the label.  -> (path_label, path_state, score)."""
        fo, n_seg, n_frames = self._align_args(frame_offsets, automata)
        a = automata
        x = _np(x, np.float32)
        if x.ndim != 2 or x.shape != (n_frames, self._in_dim(self.D)):
            raise ValueError(f"features must be [{n_frames}, {self._in_dim(self.D)}]")
        lo, lm = _np(label_offsets, np.uint32), _np(label_mixture, np.uint32)
        lab = np.empty(n_frames, np.uint32)
        st = np.empty(n_frames, np.uint32) if want_states else None
        sc = np.empty(max(n_seg, 0), np.float32)
        self._check(self._lib.gmm_score_align_labels_host(
            self._h, n_seg, _ptr(fo), _ptr(x), _ptr(a.state_offsets), _ptr(a.arc_offsets), _ptr(a.arc_target),
            _ptr(a.arc_emission), _ptr(a.arc_weight), _ptr(a.initial_state), _ptr(a.final_weight), _ptr(lo), _ptr(lm),
            _ptr(lab), _ptr(st), _ptr(sc)), "gmm_score_align_labels_host")
        return lab, st, sc

    def score_baum_welch_labels_host(self, frame_offsets, x, automata, label_offsets, label_mixture):
        """As score_baum_welch_host for arcs that carry labels -> (weights [frames, longest label list], totals);
        column j of a frame's row = posterior of label slot j of its segment."""
        fo, n_seg, n_frames = self._align_args(frame_offsets, automata)
        a = automata
        x = _np(x, np.float32)
        if x.ndim != 2 or x.shape != (n_frames, self._in_dim(self.D)):
            raise ValueError(f"features must be [{n_frames}, {self._in_dim(self.D)}]")
        lo, lm = _np(label_offsets, np.uint32), _np(label_mixture, np.uint32)
        ld_w = max(1, int(np.max(np.diff(lo)))) if n_seg else 1
        w = np.zeros((n_frames, ld_w), np.float32)
        total = np.empty(max(n_seg, 0), np.float64)
        self._check(self._lib.gmm_score_baum_welch_labels_host(
            self._h, n_seg, _ptr(fo), _ptr(x), _ptr(a.state_offsets), _ptr(a.arc_offsets), _ptr(a.arc_target),
            _ptr(a.arc_emission), _ptr(a.arc_weight), _ptr(a.initial_state), _ptr(a.final_weight), _ptr(lo), _ptr(lm),
            _ptr(w), ld_w, _ptr(total)), "gmm_score_baum_welch_labels_host")
        return w, total

    @property
    def align_ordered_count(self) -> int:
        return int(self._lib.gmm_align_ordered_count(self._h))

    # -- front end on the device (lda.flow) -------------------------------------------------------
    def set_lda(self, matrix, base_dim, length, right):
        """matrix [D, base_dim*length]; afterwards every *_host call takes base features [T, base_dim]."""
        matrix = _np(matrix, np.float32)
        if matrix.ndim != 2 or matrix.shape[1] != base_dim * length:
            raise ValueError(f"matrix must be [D, {base_dim * length}]")
        self._check(self._lib.gmm_set_lda(self._h, int(base_dim), int(length), int(right), matrix.shape[0],
                                          _ptr(matrix)), "gmm_set_lda")
        self._lda = (int(base_dim), matrix.shape[0])

    def clear_lda(self):
        self._check(self._lib.gmm_set_lda(self._h, 0, 0, 0, 0, None), "gmm_set_lda")
        self._lda = None

    def lda_device(self, frame_offsets, d_base, d_x):
        fo = _np(frame_offsets, np.uint32)
        self._check(self._lib.gmm_lda_device(self._h, fo.shape[0] - 1, _ptr(fo), _ptr(d_base), _ptr(d_x)),
                    "gmm_lda_device")

    def _in_dim(self, model_dim):
        lda = getattr(self, "_lda", None)
        return lda[0] if lda else model_dim

    # -- training E-step -------------------------------------------------------------------------
    def create_accumulators(self, mean_idx, n_means=None, viterbi=True, weight_threshold=float(np.finfo(np.float32).eps)):
        """Estimators for the uploaded float model; mean_idx [G] names the mean of every (mixture, density) slot."""
        mean_idx = _np(mean_idx, np.uint32)
        if mean_idx.shape != (self.G,):
            raise ValueError(f"mean_idx must be [{self.G}]")
        n_means = int(mean_idx.max()) + 1 if n_means is None else int(n_means)
        self._check(self._lib.gmm_accumulator_create(self._h, n_means, _ptr(mean_idx), int(bool(viterbi)),
                                                     C.c_double(float(weight_threshold))), "gmm_accumulator_create")
        self.n_means = n_means

    def reset_accumulators(self):
        self._check(self._lib.gmm_accumulator_reset(self._h), "gmm_accumulator_reset")

    def accumulate_host(self, x, mixture, weight=None, want_best=True):
        """Frames x [T, D] aligned to `mixture` [T] (0xFFFFFFFF skips) -> chosen density per frame | None."""
        x = _np(x, np.float32)
        mixture = _np(mixture, np.uint32)
        weight = None if weight is None else _np(weight, np.float64)
        if x.ndim != 2 or x.shape[1] != self._in_dim(self.D) or mixture.shape != (x.shape[0],):
            raise ValueError(f"features must be [T, {self._in_dim(self.D)}] and mixture [T]")
        best = np.empty(x.shape[0], np.uint32) if want_best else None
        self._check(self._lib.gmm_accumulate_host(self._h, x.shape[0], _ptr(x), _ptr(mixture), _ptr(weight), _ptr(best)),
                    "gmm_accumulate_host")
        return best

    def accumulate_device(self, T, d_x, d_mixture, d_weight=None, d_best=None):
        self._check(self._lib.gmm_accumulate_device(self._h, int(T), _ptr(d_x), _ptr(d_mixture), _ptr(d_weight),
                                                    _ptr(d_best)), "gmm_accumulate_device")

    @property
    def accumulator_size(self) -> int:
        return int(self._lib.gmm_accumulator_size(self._h))

    def export_accumulators(self, d_dst):
        """All estimators as one block of `accumulator_size` doubles into device memory (async on the stream)."""
        self._check(self._lib.gmm_accumulator_export_device(self._h, _ptr(d_dst)), "gmm_accumulator_export_device")

    def import_accumulators(self, d_src):
        self._check(self._lib.gmm_accumulator_import_device(self._h, _ptr(d_src)), "gmm_accumulator_import_device")

    def read_accumulators(self, n_cov):
        out = dict(mean_sum=np.empty((self.n_means, self.D)), mean_weight=np.empty(self.n_means),
                   cov_sum=np.empty((n_cov, self.D)), cov_weight=np.empty(n_cov), mix_weight=np.empty(self.G))
        self._check(self._lib.gmm_accumulator_read(self._h, _ptr(out["mean_sum"]), _ptr(out["mean_weight"]),
                                                   _ptr(out["cov_sum"]), _ptr(out["cov_weight"]),
                                                   _ptr(out["mix_weight"])), "gmm_accumulator_read")
        return out

    # -- quantized engine -------------------------------------------------------------------------
    def upload_quant(self, offsets, qmean, cov_idx, qconst, isr_scaled, scaling_squared, rule=0):
        """rule: QUANT_RULE_SIMD (f64 expression of the SIMD scorer) or QUANT_RULE_BATCH_INT (one f32 division)."""
        offsets = _np(offsets, np.uint32)
        qmean = _np(qmean, np.uint8)
        cov_idx = _np(cov_idx, np.uint32)
        qconst = _np(qconst, np.int32)
        isr_scaled = _np(isr_scaled, np.float32)
        M, G = offsets.shape[0] - 1, int(offsets[-1])
        Cn, D = isr_scaled.shape
        Dpad = qmean.shape[1]
        if qmean.shape[0] != G or cov_idx.shape != (G,) or qconst.shape != (G,):
            raise ValueError("upload_quant: table shapes are inconsistent")
        self._check(self._lib.gmm_upload_quant(self._h, M, G, D, Dpad, Cn, _ptr(offsets), _ptr(qmean), _ptr(cov_idx),
                                               _ptr(qconst), _ptr(isr_scaled), C.c_float(float(scaling_squared))),
                    "gmm_upload_quant")
        if rule:
            self._check(self._lib.gmm_quant_final_score_rule(self._h, int(rule)), "gmm_quant_final_score_rule")
        self.qM, self.qD, self.qC, self.qDpad = M, D, Cn, Dpad

    def quant_score_host(self, x, want_q=True, want_best=True, want_score=True, want_qx=False):
        x = _np(x, np.float32)
        if x.ndim != 2 or x.shape[1] != self._in_dim(self.qD):
            raise ValueError(f"features must be [T, {self._in_dim(self.qD)}]")
        T, M = x.shape[0], self.qM
        q = np.empty((T, M), np.int32) if want_q else None
        best = np.empty((T, M), np.uint32) if want_best else None
        score = np.empty((T, M), np.float32) if want_score else None
        qx = np.empty((T, self.qC, self.qDpad), np.uint8) if want_qx else None
        self._check(self._lib.gmm_quant_score_host(self._h, T, _ptr(x), _ptr(q), _ptr(best), _ptr(score), _ptr(qx), M),
                    "gmm_quant_score_host")
        return dict(q=q, best=best, score=score, qx=qx)

    def quant_score_device(self, T, d_x, d_q=None, d_best=None, d_score=None, d_qx=None, ld=None):
        ld = self.qM if ld is None else int(ld)
        self._check(self._lib.gmm_quant_score_device(self._h, int(T), _ptr(d_x), _ptr(d_q), _ptr(d_best),
                                                     _ptr(d_score), _ptr(d_qx), ld), "gmm_quant_score_device")

    def quant_score_device_static(self, T, d_x, d_q=None, d_best=None, d_score=None, ld=None):
        """gmm_quant_score_device without any allocation (scratch in a static device array; include/gmm_b200.h)."""
        ld = self.qM if ld is None else int(ld)
        self._check(self._lib.gmm_quant_score_device_static(self._h, int(T), _ptr(d_x), _ptr(d_q), _ptr(d_best),
                                                            _ptr(d_score), ld), "gmm_quant_score_device_static")
