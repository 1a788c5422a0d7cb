"""Model I/O for the GPU layout: RASR's mixture-set text file and its binary estimator
("MIXSET" accumulator) file, read straight into / written straight from the flat tables the kernels use
(`MixtureSetData`), plus the M-step that turns accumulators into a mixture set.

Reference (paths relative to /root/reference/src/Mm):
  text v2.0     MixtureSet.cc:142-214 (write/read), Mixture.cc:81-107, GaussDensity.cc:25-70 (Mean,
                DiagonalCovariance), MixtureSetTopology.cc:19-30 (GaussDensity: mean index, covariance index)
  MIXSET binary AbstractMixtureSetEstimator.cc:403-420 (header), 433-480 (read), 482-506 (write),
                VectorAccumulator.hh:75-97, GaussDensityEstimator.cc:46-63, MixtureEstimator.cc:140-170;
                little endian (Core/BinaryStream.hh:35-40)
  combine       AbstractMixtureSetEstimator.cc:149-169
  M-step        AbstractMixtureSetEstimator.cc:299-335 (estimate), MixtureEstimator.cc:38-79 (density removal),
                112-123, Mixture.cc:63-74 + Utilities.hh:44-51 (weights), GaussDensityEstimator.cc:148-156 (mean),
                200-224 (covariance), Utilities.hh:93-106,148-159

Parity: pinned against the reference binaries -- tests/golden/est_*.npz hold the text mixture set, the MIXSET
accumulator file and the estimate() output of the unmodified classes (oracle/ref_build/ref_estimator.cc); the writers
here reproduce them byte for byte (tests/test_estimator_golden.py).
"""
from __future__ import annotations

import dataclasses
import gzip
import io
import math
import struct

import numpy as np

from .gms import MixtureSetData

F64_MIN = -1.7976931348623157e308   # Core::Type<f64>::min: log-weight of a density with weight 0 (Mixture.cc:64)


def _open(path, mode):
    if str(path).endswith(".gz"):
        return gzip.open(path, mode)
    return open(path, mode)


# ---------------------------------------------------------------------------------------------------------------
# text mixture set ("#Version: 2.0")
# ---------------------------------------------------------------------------------------------------------------
def _g(v) -> str:
    """std::ostream << f32/f64 with the default precision of 6 significant digits."""
    return "%g" % float(v)


def write_mixture_set_text(ms: MixtureSetData, path) -> None:
    """MixtureSet::write(std::ostream&) (MixtureSet.cc:142-167).  Like the reference's writer this keeps only 6
    significant digits; use `write_mixture_set_text_exact` for a lossless file the reference reads the same way."""
    _write_text(ms, path, _g, _g)


def write_mixture_set_text_exact(ms: MixtureSetData, path) -> None:
    """Same grammar with round-trip digits (%.9g for f32 values, %.17g for the f64 log-weights)."""
    _write_text(ms, path, lambda v: "%.9g" % float(v), lambda v: "%.17g" % float(v))


def _write_text(ms, path, f32s, f64s):
    out = io.StringIO()
    out.write("#Version: 2.0\n#CovarianceType: DiagonalCovariance\n")
    out.write(f"{ms.D} {ms.M} {ms.n_densities} {ms.n_means} {ms.n_cov}\n")
    for m in range(ms.M):                                            # Mixture::write, Mixture.cc:81-88
        a, b = int(ms.mix_offsets[m]), int(ms.mix_offsets[m + 1])
        out.write(str(b - a))
        for k in range(a, b):
            out.write(f" {int(ms.mix_density[k])} {f64s(ms.mix_logw[k])}")
        out.write("\n")
    for d in range(ms.n_densities):                                  # GaussDensity: "mean covariance"
        out.write(f"{int(ms.dens_mean[d])} {int(ms.dens_cov[d])}\n")
    for i in range(ms.n_means):                                      # Mean::write, GaussDensity.cc:25-31
        out.write(str(ms.D) + "".join(" " + f32s(v) for v in ms.means[i]) + "\n")
    for i in range(ms.n_cov):                                        # " " << DiagonalCovariance (value weight)*
        out.write(" " + str(ms.D) + "".join(" " + f32s(v) + " 1" for v in ms.variances[i]) + "\n")
    with _open(path, "wt") as f:
        f.write(out.getvalue())


def read_mixture_set_text(path) -> MixtureSetData:
    """MixtureSet::read(std::istream&) (MixtureSet.cc:168-214): whitespace-separated tokens after two header lines.
    Version < 2.0 files store mixture weights, not log-weights (Mixture.cc:98-103)."""
    with _open(path, "rt") as f:
        line = f.readline()
        if not line.startswith("#Version:"):
            raise ValueError("not a RASR text mixture set (missing #Version line)")
        version = np.float32(float(line[10:].strip() or 0))
        if version > 2.0:
            raise ValueError(f'version "{version}" not supported')
        cov = f.readline()
        if cov[17:].strip() != "DiagonalCovariance":
            raise ValueError(f"No correct Covariance Type set: {cov[17:].strip()}")
        tok = f.read().split()
    pos = 0

    def take(n):
        nonlocal pos
        r = tok[pos:pos + n]
        if len(r) != n:
            raise ValueError("mixture set text file ends early")
        pos += n
        return r

    D, M, n_dns, n_mean, n_cov = (int(v) for v in take(5))
    offs, dens, logw = [0], [], []
    for _ in range(M):
        n = int(take(1)[0])
        pairs = take(2 * n)
        for k in range(n):
            dens.append(int(pairs[2 * k]))
            w = float(pairs[2 * k + 1])
            if version < 2.0:
                w = math.log(w) if w > 0 else F64_MIN               # Mixture::addDensity
            logw.append(w)
        offs.append(len(dens))
    dm = np.array(take(2 * n_dns), dtype=np.int64).reshape(n_dns, 2) if n_dns else np.zeros((0, 2), np.int64)
    means = np.empty((n_mean, D), np.float32)
    for i in range(n_mean):
        dim = int(take(1)[0])
        if dim != D:
            raise ValueError(f"mean {i} has dimension {dim}, expected {D}")
        means[i] = np.array(take(D), dtype=np.float32)
    var = np.empty((n_cov, D), np.float32)
    for i in range(n_cov):
        dim = int(take(1)[0])
        if dim != D:
            raise ValueError(f"covariance {i} has dimension {dim}, expected {D}")
        vw = take(2 * D)
        v = np.array(vw[0::2], dtype=np.float32).astype(np.float64)   # VarianceType value ...
        w = np.array(vw[1::2], dtype=np.float64)                      # ... times its f64 feature weight
        var[i] = (v * w).astype(np.float32)                           # DiagonalCovariance::read, GaussDensity.cc:55-70
    return MixtureSetData(D, offs, dens, logw, dm[:, 0], dm[:, 1], means, var)


# ---------------------------------------------------------------------------------------------------------------
# estimators (accumulators) and the binary MIXSET file
# ---------------------------------------------------------------------------------------------------------------
@dataclasses.dataclass
class MixtureSetEstimatorData:
    """Accumulators with the topology of a mixture set (AbstractMixtureSetEstimator's state)."""
    D: int
    mix_offsets: np.ndarray   # u32 [M+1]
    mix_density: np.ndarray   # u32 [G]   density estimator of slot (mixture, k)
    mix_weight: np.ndarray    # f64 [G]   AbstractMixtureEstimator::weights_
    dens_mean: np.ndarray     # u32 [nDens]
    dens_cov: np.ndarray      # u32 [nDens]
    mean_sum: np.ndarray      # f64 [nMeans, D]
    mean_weight: np.ndarray   # f64 [nMeans]
    cov_sum: np.ndarray       # f64 [nCov, D]   sum of w * x^2
    cov_weight: np.ndarray    # f64 [nCov]

    def __post_init__(self):
        u, f = np.uint32, np.float64
        self.mix_offsets = np.ascontiguousarray(self.mix_offsets, u)
        self.mix_density = np.ascontiguousarray(self.mix_density, u)
        self.mix_weight = np.ascontiguousarray(self.mix_weight, f)
        self.dens_mean = np.ascontiguousarray(self.dens_mean, u)
        self.dens_cov = np.ascontiguousarray(self.dens_cov, u)
        self.mean_sum = np.ascontiguousarray(self.mean_sum, f).reshape(-1, self.D)
        self.mean_weight = np.ascontiguousarray(self.mean_weight, f)
        self.cov_sum = np.ascontiguousarray(self.cov_sum, f).reshape(-1, self.D)
        self.cov_weight = np.ascontiguousarray(self.cov_weight, f)

    @property
    def M(self):
        return self.mix_offsets.shape[0] - 1

    @classmethod
    def from_accumulators(cls, ms: MixtureSetData, acc: dict) -> "MixtureSetEstimatorData":
        """The dict of `Engine.read_accumulators` / `oracle.accumulate` on the topology of `ms`."""
        return cls(ms.D, ms.mix_offsets.copy(), ms.mix_density.copy(), acc["mix_weight"], ms.dens_mean.copy(),
                   ms.dens_cov.copy(), acc["mean_sum"], acc["mean_weight"], acc["cov_sum"], acc["cov_weight"])

    def equal_topology(self, o: "MixtureSetEstimatorData") -> bool:
        return (self.D == o.D and np.array_equal(self.mix_offsets, o.mix_offsets)
                and np.array_equal(self.mix_density, o.mix_density) and np.array_equal(self.dens_mean, o.dens_mean)
                and np.array_equal(self.dens_cov, o.dens_cov))

    def accumulate(self, o: "MixtureSetEstimatorData") -> bool:
        """Combine two accumulator sets (AbstractMixtureSetEstimator.cc:149-169); False if the topologies differ."""
        if not self.equal_topology(o):
            return False
        self.mix_weight += o.mix_weight
        self.mean_sum += o.mean_sum
        self.mean_weight += o.mean_weight
        self.cov_sum += o.cov_sum
        self.cov_weight += o.cov_weight
        return True

    # -- first-appearance numbering of MixtureSetEstimatorIndexMap (AbstractMixtureSetEstimator.cc:804-817) -----
    def _index_maps(self, slots):
        dens_map, mean_map, cov_map = {}, {}, {}
        for g in slots:
            d = int(self.mix_density[g])
            mean_map.setdefault(int(self.dens_mean[d]), len(mean_map))
            cov_map.setdefault(int(self.dens_cov[d]), len(cov_map))
            dens_map.setdefault(d, len(dens_map))
        return dens_map, mean_map, cov_map

    # -- M-step ------------------------------------------------------------------------------------------
    def estimate(self, *, min_observation_weight=5.0, min_relative_weight=0.0, min_variance=0.0,
                 normalize_mixture_weights=True, allow_zero_weights=False) -> MixtureSetData:
        """AbstractMixtureSetEstimator::estimate (.cc:306-335) with the reference's parameter defaults."""
        M = self.M
        if not allow_zero_weights:
            for m in range(M):
                if self.mix_weight[self.mix_offsets[m]:self.mix_offsets[m + 1]].sum() == 0:
                    raise ValueError(f"Mixture {m} has zero weight.")
        # covariance -> set of its mean estimators, BEFORE densities are removed (.cc:311-314)
        all_slots = range(int(self.mix_offsets[-1]))
        cov_means = {}
        for g in all_slots:
            d = int(self.mix_density[g])
            cov_means.setdefault(int(self.dens_cov[d]), set()).add(int(self.dens_mean[d]))
        # removeDensitiesWithLowWeight per mixture (MixtureEstimator.cc:64-79): the heaviest density always stays
        kept = []
        for m in range(M):
            a, b = int(self.mix_offsets[m]), int(self.mix_offsets[m + 1])
            if a == b:
                kept.append([])
                continue
            w = self.mix_weight[a:b]
            heaviest = int(np.argmax(w))                           # first maximum (MixtureEstimator.cc:38-47)
            min_w = max(float(min_observation_weight), float(w.sum()) * float(min_relative_weight))
            kept.append([a + k for k in range(b - a) if w[k] >= min_w or k == heaviest])
        slots = [g for ks in kept for g in ks]
        dens_map, mean_map, cov_map = self._index_maps(slots)
        offs, dens, logw = [0], [], []
        for ks in kept:
            lw = np.array([math.log(self.mix_weight[g]) if self.mix_weight[g] > 0 else F64_MIN for g in ks])
            if normalize_mixture_weights and len(ks):               # Mixture::normalizeWeights + logExpNorm
                i_max = int(np.argmax(lw))
                rest = sum(math.exp(lw[i] - lw[i_max]) for i in range(len(ks)) if i != i_max)
                lw = lw - (math.log1p(rest) + lw[i_max])
            dens += [dens_map[int(self.mix_density[g])] for g in ks]
            logw += list(lw)
            offs.append(len(dens))
        n_d, n_m, n_c = len(dens_map), len(mean_map), len(cov_map)
        dens_mean, dens_cov = np.zeros(n_d, np.uint32), np.zeros(n_d, np.uint32)
        for d, i in dens_map.items():
            dens_mean[i] = mean_map[int(self.dens_mean[d])]
            dens_cov[i] = cov_map[int(self.dens_cov[d])]
        means = np.zeros((n_m, self.D), np.float32)
        for mi, i in mean_map.items():                               # MeanEstimator::estimate
            if self.mean_weight[mi] != 0:
                means[i] = (self.mean_sum[mi] / self.mean_weight[mi]).astype(np.float32)
        var = np.ones((n_c, self.D), np.float32)                     # weight 0: DiagonalCovariance(dim) = ones
        for ci, i in cov_map.items():                                # CovarianceEstimator::estimate
            if self.cov_weight[ci] == 0:
                continue
            mean_sq = np.zeros(self.D)
            for mi in sorted(cov_means[ci]):
                if self.mean_weight[mi] > 0:
                    mean_sq = mean_sq + self.mean_sum[mi] * self.mean_sum[mi] / self.mean_weight[mi]
            v = ((self.cov_sum[ci] - mean_sq) / self.cov_weight[ci]).astype(np.float32)
            if min_variance != 0:
                v = np.maximum(v, np.float32(min_variance))
            var[i] = v
        return MixtureSetData(self.D, offs, dens, logw, dens_mean, dens_cov, means, var)


MIXSET_MAGIC = b"MIXSET\0\0"   # os.write(magic().c_str(), 8): "MIXSET", its NUL and one more byte


def write_mixset(est: MixtureSetEstimatorData, path, version: int = 2) -> None:
    """AbstractMixtureSetEstimator::write (.cc:482-506), estimators numbered by first appearance."""
    if version < 1:
        raise ValueError("only the f64-weight layout (version >= 1) is written")
    dens_map, mean_map, cov_map = est._index_maps(range(int(est.mix_offsets[-1])))
    out = io.BytesIO()
    out.write(MIXSET_MAGIC)
    out.write(struct.pack("<II", version, est.D))

    def accumulators(index_map, sums, weights):
        out.write(struct.pack("<I", len(index_map)))
        for src in index_map:                                        # dict keeps first-appearance order
            out.write(struct.pack("<I", est.D))                      # VectorAccumulator::write: size, sums, weight
            out.write(np.ascontiguousarray(sums[src], "<f8").tobytes())
            out.write(struct.pack("<d", float(weights[src])))

    accumulators(mean_map, est.mean_sum, est.mean_weight)
    accumulators(cov_map, est.cov_sum, est.cov_weight)
    out.write(struct.pack("<I", len(dens_map)))
    for d in dens_map:
        out.write(struct.pack("<II", mean_map[int(est.dens_mean[d])], cov_map[int(est.dens_cov[d])]))
    out.write(struct.pack("<I", est.M))
    for m in range(est.M):
        a, b = int(est.mix_offsets[m]), int(est.mix_offsets[m + 1])
        out.write(struct.pack("<I", b - a))
        for g in range(a, b):
            out.write(struct.pack("<Id", dens_map[int(est.mix_density[g])], float(est.mix_weight[g])))
    with _open(path, "wb") as f:
        f.write(out.getvalue())


def read_mixset(path) -> MixtureSetEstimatorData:
    """AbstractMixtureSetEstimator::read (.cc:433-480); version 0 files carry u32 counts instead of f64 weights."""
    with _open(path, "rb") as f:
        buf = f.read()
    magic = buf[:8].split(b"\0")[0]
    if magic != b"MIXSET":
        raise ValueError(f'Mixture set estimator file with magic "{magic.decode(errors="replace")}" could not be '
                         'read, "MIXSET" expected.')
    pos = 8

    def take(fmt):
        nonlocal pos
        v = struct.unpack_from("<" + fmt, buf, pos)
        pos += struct.calcsize("<" + fmt)
        return v if len(v) > 1 else v[0]

    version, D = take("II")

    def accumulators():
        n = take("I")
        sums, weights = np.zeros((n, D)), np.zeros(n)
        for i in range(n):
            size = take("I")
            if size != D:
                raise ValueError(f"accumulator of size {size} in a mixture set estimator of dimension {D}")
            nonlocal pos
            sums[i] = np.frombuffer(buf, "<f8", D, pos)
            pos += 8 * D
            weights[i] = take("d") if version > 0 else float(take("I"))
        return sums, weights

    mean_sum, mean_w = accumulators()
    cov_sum, cov_w = accumulators()
    n_dens = take("I")
    dm = np.array([take("II") for _ in range(n_dens)], np.uint32).reshape(n_dens, 2)
    M = take("I")
    offs, dens, w = [0], [], []
    for _ in range(M):
        for _ in range(take("I")):
            dens.append(take("I"))
            w.append(take("d") if version > 0 else float(take("I")))
        offs.append(len(dens))
    if pos != len(buf):
        raise ValueError("error reading mixture set estimator (trailing bytes)")
    return MixtureSetEstimatorData(D, offs, dens, w, dm[:, 0], dm[:, 1], mean_sum, mean_w, cov_sum, cov_w)
