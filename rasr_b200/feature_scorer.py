"""Host-side mirror of the reference's scorer plugin surface over the C-ABI (tests / bench convenience).

The production host side is the C++ class in rasr/GpuFeatureScorer.{hh,cc}; this module states the same
interface in Python so that the parity tests read like calls into the reference:

* ``FeatureScorer`` interface and buffered protocol   /root/reference/src/Mm/FeatureScorer.hh:52-134
* ``AssigningFeatureScorer``                          /root/reference/src/Mm/AssigningFeatureScorer.hh:60-121
* ring-buffer state machine                           /root/reference/src/Mm/BatchFeatureScorer.cc:40-105
* whole-segment ("never filled") precedent            /root/reference/src/Nn/TrainerFeatureScorer.cc:160-230
* factory names -> scorer                             /root/reference/src/Mm/Module.cc:75-108,301-309

Method names follow the reference (camelCase).  Contract violations raise ``AssertionError`` where the
reference uses ``require()``; configuration errors raise ``ValueError`` where it uses ``criticalError()``.
Frames are scored in batches on the GPU: the first request for a frame that has no score yet sends every
pending frame through one kernel launch; ``score(e)`` afterwards is an array read.
"""
from __future__ import annotations

import collections

import numpy as np

from . import capi, tables
from .gms import MixtureSetData

U32_MAX = 0xFFFFFFFF


class _ScoreBlock:
    """Scores of one GPU batch, kept alive by the context scorers that point into it (the reference hands
    out Core::Ref<const ContextScorer> that may outlive the next getScorer call, AlignmentNode.cc:281-306)."""
    __slots__ = ("scores", "best")

    def __init__(self, scores, best):
        self.scores, self.best = scores, best


class ContextScorer:
    """Mm::FeatureScorer::ContextScorer / AssigningContextScorer: one frame's emission scores."""
    __slots__ = ("_parent", "_frame", "_block", "_row")

    def __init__(self, parent, frame):
        self._parent, self._frame, self._block, self._row = parent, frame, None, -1

    def _resolve(self):
        if self._block is None:
            self._parent._score_pending()
            assert self._block is not None
        return self._block

    def nEmissions(self) -> int:
        return self._parent.nMixtures()

    def score(self, e: int) -> float:
        assert 0 <= e < self._parent.nMixtures()
        return float(self._resolve().scores[self._row, e])

    def bestDensity(self, e: int) -> int:
        b = self._resolve().best
        if b is None:
            raise ValueError("scorer was created without density assignment (assigning=False)")
        return int(b[self._row, e])

    def scoreOfDensity(self, e: int, dns_in_mix: int) -> float:
        """score(e, dnsInMix) (AssigningFeatureScorer.hh:46,122-134): a criticalError in both reference diagonal scorers
        (neither overrides calculateScore(cs, e, dns)), a ValueError here, as in the C++ plugin."""
        raise ValueError("This feature scorer does not support the calculation of scores given a density in the mixture")

    def getDensityPosteriorProbabilities(self, e: int) -> np.ndarray:
        """Density posteriors exp(score - s_k) of the sum scorer (GaussDiagonalMaximumFeatureScorer.cc:291-298, consumer
        AbstractMixtureSetEstimator.cc:386-395); the other scorers answer like the reference (criticalError)."""
        assert 0 <= e < self._parent.nMixtures()
        if self._parent._type != "diagonal-sum":
            raise ValueError("This feature scorer does not support the calculation of density posterior probabilities")
        self._resolve()
        k = self._parent.densitiesInMixture(e)
        widest = int(np.max(np.diff(self._parent._offsets)))  # the call's row pitch covers the model's largest mixture
        post, _ = self._parent._engine.mixture_posteriors_host(self._frame[None, :], np.array([e], np.uint32), max(1, widest))
        return post[0, :k]

    def scores(self) -> np.ndarray:
        """All M scores of this frame (eager consumers, Speech/FeatureScorerNode.cc:95-155)."""
        return self._resolve().scores[self._row, : self._parent.nMixtures()]


class GpuFeatureScorer:
    """Buffered GPU feature scorer.  ``scorer_type`` is the value of the reference's ``feature-scorer-type``:

    ``diagonal-maximum``       float, maximum approximation         (GaussDiagonalMaximumFeatureScorer)
    ``diagonal-sum``           float, full sum / log-sum-exp        (GaussDiagonalSumFeatureScorer)
    ``SIMD-diagonal-maximum``  u8 quantized, bit-exact integer path (SimdGaussDiagonalMaximumFeatureScorer)
    ``batch-diagonal-maximum-int``  the same u8 kernel with the constants and the final scale of
                               BatchIntFeatureScorer (BatchFeatureScorer.cc:319-376,468); pooled covariance only
    ``batch-diagonal-maximum-fast`` BatchUnrolledIntFeatureScorer (:537-604): the batch-int numbers, at most 48 padded
                               components
    ``batch-diagonal-maximum-float`` BatchFloatFeatureScorer (:138-234): the float maximum scorer over whitened
                               operands, pooled covariance only (within the float tolerance)

    ``buffer_size`` mirrors the reference's ``buffer-size`` parameter; 0 selects segment mode (the buffer
    never reports filled, everything is scored in one launch when the first score is requested).
    """

    SCORER_TYPES = ("diagonal-maximum", "diagonal-sum", "SIMD-diagonal-maximum", "batch-diagonal-maximum-int",
                    "batch-diagonal-maximum-fast", "batch-diagonal-maximum-float")
    QUANTIZED_TYPES = ("SIMD-diagonal-maximum", "batch-diagonal-maximum-int", "batch-diagonal-maximum-fast")
    BATCH_INT_TYPES = ("batch-diagonal-maximum-int", "batch-diagonal-maximum-fast")

    def __init__(self, mixture_set: MixtureSetData, scorer_type: str = "diagonal-maximum", *, buffer_size: int = 0,
                 device: int = 0, assigning: bool = True, kernel: int = capi.KERNEL_AUTO,
                 mixture_weight_scale: float = 1.0, gaussian_scale: float = 1.0, quant_tables: dict | None = None,
                 float_tables: dict | None = None, keep_frames: bool = False):
        if scorer_type not in self.SCORER_TYPES:
            raise ValueError(f"unknown feature-scorer-type '{scorer_type}'")  # Module.cc:303-307
        if buffer_size < 0:
            raise ValueError("buffer-size must be >= 0")
        if scorer_type == "batch-diagonal-maximum-float":
            if mixture_set.n_cov != 1:
                raise ValueError("feature scorer supports only globally pooled covariance")  # BatchFeatureScorer.cc:146-148
            scorer_type = "diagonal-maximum"
        if scorer_type == "batch-diagonal-maximum-fast" and tables.simd_padded_dimension(mixture_set.D) > 48:
            raise ValueError("This feature scorer supports only features with max. 48 components")  # :550-553
        self._type = scorer_type
        self._M, self._D = mixture_set.M, mixture_set.D
        self._offsets = mixture_set.mix_offsets.copy()
        self._buffer_size = int(buffer_size)
        self._assigning = assigning
        self._keep_frames = keep_frames   # contexts keep their frame after scoring (rasr_b200.aligner.Aligner reads it)
        self._engine = capi.Engine(device)
        if scorer_type in self.QUANTIZED_TYPES:
            if quant_tables is not None:
                t = quant_tables
            elif scorer_type == "SIMD-diagonal-maximum":
                t = tables.quant_tables(mixture_set)
            else:
                t = tables.batchint_tables(mixture_set)
            self._engine.upload_quant(t["offsets"], t["qmean"], t["cov_idx"], t["qconst"], t["isr_scaled"],
                                      t["scaling_squared"],
                                      rule=capi.QUANT_RULE_BATCH_INT if scorer_type in self.BATCH_INT_TYPES
                                      else capi.QUANT_RULE_SIMD)
            self.scaling_squared = np.float32(t["scaling_squared"])
        else:
            t = float_tables if float_tables is not None else tables.float_tables(
                mixture_set, mixture_weight_scale=mixture_weight_scale, gaussian_scale=gaussian_scale)
            mode = capi.MODE_SUM if scorer_type == "diagonal-sum" else capi.MODE_MAX
            self._engine.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"],
                                      mode=mode, kernel=kernel)
        self._fifo = collections.deque()  # ContextScorer objects in time order, oldest first
        self._pending = []  # ContextScorers without a score block yet (time order)
        self._segment = []  # segment mode (buffer-size 0): every frame since reset(), addressable by time
        self.n_batches = 0

    # -- FeatureScorer interface -------------------------------------------------------------------
    def nMixtures(self) -> int:
        return self._M

    def dimension(self) -> int:
        return self._D

    def nDensities(self) -> int:
        return int(self._offsets[-1])

    def densitiesInMixture(self, m: int) -> int:
        return int(self._offsets[m + 1] - self._offsets[m])

    def isBuffered(self) -> bool:
        return True

    def bufferSize(self) -> int:
        return self._buffer_size if self._buffer_size > 0 else U32_MAX

    def bufferFilled(self) -> bool:
        return self._buffer_size > 0 and len(self._fifo) >= self._buffer_size - 1

    def bufferEmpty(self) -> bool:
        return len(self._fifo) == 0

    def reset(self) -> None:
        self._fifo.clear()
        self._pending = []
        self._segment = []

    def hasTimeIndexedCache(self) -> bool:
        """FeatureScorer.hh:157-159; true in segment mode, where every frame of the segment is kept."""
        return self._buffer_size == 0

    def getTimeIndexedScorer(self, time: int) -> "ContextScorer":
        """FeatureScorer.hh:160-163, used by Speech/SegmentwiseAlignmentGenerator.cc:90-156."""
        if not self.hasTimeIndexedCache():
            raise RuntimeError("time-indexed scores not available")
        assert 0 <= time < len(self._segment)
        return self._segment[time]

    def _push(self, f):
        f = np.asarray(f, dtype=np.float32)
        if f.shape != (self._D,):
            raise AssertionError(f"feature dimension {f.shape} != {self._D}")  # require(dimension == size)
        cs = ContextScorer(self, f)
        self._fifo.append(cs)
        self._pending.append(cs)
        if self._buffer_size == 0:
            self._segment.append(cs)

    def addFeature(self, f) -> None:
        assert not self.bufferFilled(), "addFeature requires !bufferFilled()"
        self._push(f)

    def getScorer(self, f) -> ContextScorer:
        """Add f, return the scorer of the OLDEST buffered frame (BatchFeatureScorer.cc:77-87)."""
        self._push(f)
        return self._fifo.popleft()

    getAssigningScorer = getScorer

    def flush(self) -> ContextScorer:
        assert not self.bufferEmpty(), "flush requires !bufferEmpty()"
        return self._fifo.popleft()

    # -- batch scoring ----------------------------------------------------------------------------
    def _score_pending(self) -> None:
        if not self._pending:
            return
        x = np.stack([c._frame for c in self._pending])
        scores, best = self.scoreBatch(x)
        block = _ScoreBlock(scores, best)
        keep = self._type == "diagonal-sum" or self._keep_frames  # density posteriors / the aligner need the frame later
        for i, c in enumerate(self._pending):
            c._block, c._row, c._frame = block, i, (c._frame if keep else None)
        self._pending = []
        self.n_batches += 1

    def scoreBatch(self, x: np.ndarray):
        """Eager scoring of a [T, D] block: (scores [T, M] f32, best [T, M] u32 or None)."""
        if self._type in self.QUANTIZED_TYPES:
            r = self._engine.quant_score_host(x, want_q=False, want_best=self._assigning, want_score=True)
            return r["score"], r["best"]
        return self._engine.score_host(x, want_best=self._assigning)

    @property
    def engine(self) -> capi.Engine:
        return self._engine


def createFeatureScorer(config: dict, mixture_set: MixtureSetData) -> GpuFeatureScorer:
    """Mm::Module_::createFeatureScorer (Module.cc:301-309): pick the scorer by ``feature-scorer-type``."""
    return GpuFeatureScorer(mixture_set, config.get("feature-scorer-type", "diagonal-maximum"),
                            buffer_size=int(config.get("buffer-size", 0)), device=int(config.get("device", 0)),
                            mixture_weight_scale=float(config.get("mixture-weight-scale", 1.0)),
                            gaussian_scale=float(config.get("gaussian-scale", 1.0)))
