"""Raw mixture-set container ("GMS1" file) shared by the product, the tests and the oracle harness.

It mirrors the index-linked tables of the reference's ``Mm::MixtureSet``
(/root/reference/src/Mm/MixtureSet.hh:47-50,144-204): mixtures -> densities -> (mean, covariance), where
densities, means and covariances may be shared, plus per-(mixture, density) f64 log-weights
(/root/reference/src/Mm/Mixture.hh:33-35).  Nothing here computes scores.

File layout (little endian)::

    "GMS1"  u32 D  u32 M  u32 nDensities  u32 nMeans  u32 nCovariances
    u32 mixOffsets[M+1]          prefix sums of densities-per-mixture
    u32 mixDensity[total]        density index of the k-th density of each mixture
    f64 mixLogWeight[total]      log mixture weight (may be -DBL_MAX for weight 0, Mixture.cc:63-66)
    u32 densMean[nDensities]  u32 densCov[nDensities]
    f32 means[nMeans x D]     f32 variances[nCovariances x D]   (diagonal, sigma^2)
"""
from __future__ import annotations

import dataclasses
import struct

import numpy as np

MAGIC = b"GMS1"


@dataclasses.dataclass
class MixtureSetData:
    D: int
    mix_offsets: np.ndarray  # u32 [M+1]
    mix_density: np.ndarray  # u32 [total]
    mix_logw: np.ndarray  # f64 [total]
    dens_mean: np.ndarray  # u32 [nDens]
    dens_cov: np.ndarray  # u32 [nDens]
    means: np.ndarray  # f32 [nMeans, D]
    variances: np.ndarray  # f32 [nCov, D]

    def __post_init__(self):
        self.mix_offsets = np.ascontiguousarray(self.mix_offsets, dtype=np.uint32)
        self.mix_density = np.ascontiguousarray(self.mix_density, dtype=np.uint32)
        self.mix_logw = np.ascontiguousarray(self.mix_logw, dtype=np.float64)
        self.dens_mean = np.ascontiguousarray(self.dens_mean, dtype=np.uint32)
        self.dens_cov = np.ascontiguousarray(self.dens_cov, dtype=np.uint32)
        self.means = np.ascontiguousarray(self.means, dtype=np.float32).reshape(-1, self.D)
        self.variances = np.ascontiguousarray(self.variances, dtype=np.float32).reshape(-1, self.D)
        self.validate()

    # -- shape helpers ------------------------------------------------------------------------
    @property
    def M(self) -> int:
        return int(self.mix_offsets.shape[0] - 1)

    @property
    def G(self) -> int:
        """Total Gaussians = sum over mixtures of nDensities (SURVEY.md section 8 notation)."""
        return int(self.mix_offsets[-1])

    @property
    def n_densities(self) -> int:
        return int(self.dens_mean.shape[0])

    @property
    def n_means(self) -> int:
        return int(self.means.shape[0])

    @property
    def n_cov(self) -> int:
        return int(self.variances.shape[0])

    def validate(self) -> None:
        if self.mix_offsets.ndim != 1 or self.mix_offsets.shape[0] < 1 or self.mix_offsets[0] != 0:
            raise ValueError("mix_offsets must be a prefix-sum array starting at 0")
        if np.any(np.diff(self.mix_offsets.astype(np.int64)) < 0):
            raise ValueError("mix_offsets must be non-decreasing")
        if self.mix_density.shape[0] != self.G or self.mix_logw.shape[0] != self.G:
            raise ValueError("mix_density / mix_logw length must equal mix_offsets[-1]")
        if self.dens_cov.shape[0] != self.n_densities:
            raise ValueError("dens_mean / dens_cov length mismatch")
        if self.G and int(self.mix_density.max()) >= self.n_densities:
            raise ValueError("density index out of range")
        if self.n_densities and (int(self.dens_mean.max()) >= self.n_means or int(self.dens_cov.max()) >= self.n_cov):
            raise ValueError("mean / covariance index out of range")
        if not np.all(self.variances >= 0):
            raise ValueError("variances must not be negative")

    def check_scorable(self) -> None:
        """A SCORER needs strictly positive variances (reference: require(checkDiagonal(diagonal)) in
        CovarianceFeatureScorerElement.cc:23,41-44).  A mixture set as such may hold zeros: the reference's estimate()
        returns them for a covariance seen with one observation when minimum-variance is 0
        (AbstractMixtureSetEstimator.cc:299-335), and such a set can be written and read."""
        if np.any(self.variances <= 0):
            raise ValueError("variances must be strictly positive (raise minimum-variance in the M-step)")

    # -- file io ------------------------------------------------------------------------------
    def save(self, path: str) -> None:
        with open(path, "wb") as f:
            f.write(MAGIC)
            f.write(struct.pack("<5I", self.D, self.M, self.n_densities, self.n_means, self.n_cov))
            for a in (self.mix_offsets, self.mix_density, self.mix_logw, self.dens_mean, self.dens_cov,
                      self.means, self.variances):
                f.write(a.tobytes())

    @classmethod
    def load(cls, path: str) -> "MixtureSetData":
        with open(path, "rb") as f:
            buf = f.read()
        if buf[:4] != MAGIC:
            raise ValueError(f"{path}: not a GMS1 file")
        D, M, nD, nM, nC = struct.unpack_from("<5I", buf, 4)
        pos = 24

        def take(dtype, n):
            nonlocal pos
            a = np.frombuffer(buf, dtype=dtype, count=n, offset=pos).copy()
            pos += a.nbytes
            return a

        offs = take(np.uint32, M + 1)
        total = int(offs[-1])
        return cls(D, offs, take(np.uint32, total), take(np.float64, total), take(np.uint32, nD),
                   take(np.uint32, nD), take(np.float32, nM * D), take(np.float32, nC * D))
