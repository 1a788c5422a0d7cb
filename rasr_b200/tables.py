"""Host-side table preparation: MixtureSetData -> the flat per-slot tables the C-ABI uploads.

This is the packing step of the reference's scorer constructors, done once per model on the host:

* float tables   GaussDiagonalMaximumFeatureScorer::init   /root/reference/src/Mm/GaussDiagonalMaximumFeatureScorer.cc:64-86
                 CovarianceFeatureScorerElement            .../CovarianceFeatureScorerElement.cc:20-51
                 MixtureFeatureScorerElement               .../MixtureFeatureScorerElement.cc:21-33
                 gaussLogNormFactor                        .../Utilities.hh:52-76
* quantized      SimdGaussDiagonalMaximumFeatureScorer::init / getScaling / quantizationScalingFactor /
                 buildMixtureTable                         .../SimdFeatureScorer.cc:64-133
                 createDensityElement / multiplyAndQuantize .../IntelOptimization.cc:39-67
                 quantize<f32,u8>                          .../Utilities.hh:179-191

Everything is evaluated in IEEE f32/f64 with the reference's operation order.  The reference's own build
(-ffast-math) stores rsqrtss+Newton inverse standard deviations, which can differ from these
in the last bit; inside RASR the scorer class (rasr/GpuFeatureScorer.cc) therefore takes the tables from the
reference's own element classes, and the bit-exact tests feed the engine the tables of the compiled reference.
No scores are computed here.
"""
from __future__ import annotations

import numpy as np

from .gms import MixtureSetData

F32 = np.float32
INT_INDEFINITE = np.int32(-2 ** 31)  # cvttss2si result for NaN / out-of-range input


def flatten(ms: MixtureSetData):
    """Per-slot (mixture, density) views: mean [G, D], cov_idx [G]; tying is resolved by copying, as
    BatchFeatureScorerBase::initialize does (BatchFeatureScorer.cc:62-75)."""
    dns = ms.mix_density
    mean = np.ascontiguousarray(ms.means[ms.dens_mean[dns]], dtype=F32)
    cov_idx = np.ascontiguousarray(ms.dens_cov[dns], dtype=np.uint32)
    return mean, cov_idx


def mean_index(ms: MixtureSetData) -> np.ndarray:
    """Mean of every (mixture, density) slot, for the E-step's estimators (gmm_accumulator_create)."""
    return np.ascontiguousarray(ms.dens_mean[ms.mix_density], dtype=np.uint32)


def _log_norm(variances: np.ndarray) -> np.ndarray:
    D = variances.shape[1]
    # sequential f64 sum over d, as the reference's std::accumulate
    s = np.add.accumulate(np.log(np.abs(variances.astype(np.float64))), axis=1)[:, -1]
    return (np.float64(D) * np.log(np.float64(2) * np.pi) + s).astype(F32)


def float_tables(ms: MixtureSetData, *, mixture_weight_scale: float = 1.0, gaussian_scale: float = 1.0):
    """-> dict(offsets, mean [G,D], cov_idx [G], isr [C,D], log_norm [C], w2 [G])."""
    ms.check_scorable()
    mean, cov_idx = flatten(ms)
    factor = F32(np.sqrt(np.float64(gaussian_scale)))
    with np.errstate(over="ignore"):
        isr = (F32(1) / np.sqrt(ms.variances.astype(F32))).astype(F32)
        isr = (isr * factor).astype(F32)
        log_norm = (_log_norm(ms.variances) * F32(factor * factor)).astype(F32)
        w2 = ((-2.0 * ms.mix_logw).astype(F32) * F32(mixture_weight_scale)).astype(F32)
    return dict(offsets=ms.mix_offsets.copy(), mean=mean, cov_idx=cov_idx, isr=isr, log_norm=log_norm, w2=w2)


def _round_half_away(v: np.ndarray) -> np.ndarray:
    with np.errstate(invalid="ignore"):
        a = np.abs(v)
        f = np.floor(a)
        r = f + ((a - f) >= F32(0.5)).astype(F32)
        return np.copysign(r, v).astype(F32)


def _cvtt_s32(v: np.ndarray) -> np.ndarray:
    """(int) of an f32 as x86 cvttss2si does it: truncation, 0x80000000 for NaN / out of range."""
    v = np.asarray(v, dtype=F32)
    ok = np.isfinite(v) & (np.abs(v) < F32(2147483648.0))
    out = np.full(v.shape, INT_INDEFINITE, dtype=np.int32)
    out[ok] = np.trunc(v[ok]).astype(np.int32)
    return out


def quantize_u8(v: np.ndarray) -> np.ndarray:
    """quantize<f32,u8>: clip((int)round(v) + 128, 0, 255) with wrapping int add."""
    r = _cvtt_s32(_round_half_away(np.asarray(v, dtype=F32)))
    s = (r.astype(np.int64) + 128)
    s = np.where(s >= 2 ** 31, s - 2 ** 32, s)
    s = np.where(r == INT_INDEFINITE, np.int64(INT_INDEFINITE) + 128, s)
    return np.clip(s, 0, 255).astype(np.uint8)


def simd_padded_dimension(D: int) -> int:
    return ((D + 15) // 16) * 16


def quant_tables(ms: MixtureSetData):
    """-> dict(offsets, qmean [G,Dpad] u8, cov_idx [G], qconst [G] s32, isr_scaled [C,D], scaling_squared)."""
    ms.check_scorable()
    mean, cov_idx = flatten(ms)
    D, Dpad = ms.D, simd_padded_dimension(ms.D)
    with np.errstate(over="ignore", invalid="ignore"):
        isr = (F32(1) / np.sqrt(ms.variances.astype(F32))).astype(F32)
        log_norm = _log_norm(ms.variances)
        # getScaling(): over all densities (not slots) of mean * isr
        dm = (ms.means[ms.dens_mean] * isr[ms.dens_cov]).astype(F32)
        if dm.size:
            mn, mx = F32(dm.min()), F32(dm.max())
        else:
            mn, mx = F32(np.finfo(F32).max), F32(-np.finfo(F32).max)
        interval = F32(2) * max(F32(abs(mn)), F32(abs(mx)))
        scaling = F32(np.float64(F32(255)) / (1.25 * np.float64(interval)))
        s2 = F32(scaling * scaling)
        isr_s = (isr * scaling).astype(F32)
        log_norm_s = (log_norm * F32(scaling * scaling)).astype(F32)
        qmean = np.zeros((ms.G, Dpad), np.uint8)
        qmean[:, :D] = quantize_u8((mean * isr_s[cov_idx]).astype(F32))
        w = (np.float64(F32(s2 * F32(-2))) * ms.mix_logw).astype(F32)
        qconst = _cvtt_s32((w + log_norm_s[cov_idx]).astype(F32))
    return dict(offsets=ms.mix_offsets.copy(), qmean=qmean, cov_idx=cov_idx, qconst=qconst, isr_scaled=isr_s,
                scaling_squared=s2)


def batchint_tables(ms: MixtureSetData):
    """Tables of BatchIntFeatureScorer::init (`batch-diagonal-maximum-int`, src/Mm/BatchFeatureScorer.cc:319-376):
    the u8 grid of the SIMD scorer with its own constants, c = (s32)(L * scale^2 - 2 scale^2 * logWeight) evaluated in
    double, and the final score q / (2 scale^2) (:468).  Pooled covariance only, like the reference (:340-342).
    Same dict as quant_tables: the device kernel is the same; scaling_squared is scale^2 (the kernel divides by
    2 * scaling_squared)."""
    ms.check_scorable()
    if ms.n_cov != 1:
        raise ValueError("feature scorer supports only globally pooled variance")
    mean, cov_idx = flatten(ms)
    D, Dpad = ms.D, simd_padded_dimension(ms.D)
    with np.errstate(over="ignore", invalid="ignore"):
        isr = (F32(1) / np.sqrt(ms.variances.astype(F32))).astype(F32)
        log_norm = _log_norm(ms.variances)
        dm = (ms.means[ms.dens_mean] * isr[0]).astype(F32)
        if dm.size:
            mn, mx = F32(dm.min()), F32(dm.max())
        else:
            mn, mx = F32(np.finfo(F32).max), F32(-np.finfo(F32).max)
        interval = F32(2) * max(F32(abs(mn)), F32(abs(mx)))
        scale = F32(np.float64(F32(255)) / (1.25 * np.float64(interval)))
        s2 = F32(scale * scale)
        scale2 = F32(2.0 * np.float64(s2))
        isr_s = (isr * scale).astype(F32)
        log_norm_factor = F32(log_norm[0] * s2)
        qmean = np.zeros((ms.G, Dpad), np.uint8)
        qmean[:, :D] = quantize_u8((mean * isr_s[0]).astype(F32))
        c = np.float64(log_norm_factor) - np.float64(scale2) * ms.mix_logw.astype(np.float64)
        ok = np.isfinite(c) & (np.abs(c) < 2147483648.0)
        qconst = np.full(c.shape, INT_INDEFINITE, np.int32)  # cvttsd2si out of range
        qconst[ok] = np.trunc(c[ok]).astype(np.int32)
    return dict(offsets=ms.mix_offsets.copy(), qmean=qmean, cov_idx=cov_idx, qconst=qconst, isr_scaled=isr_s,
                scaling_squared=F32(scale2 * F32(0.5)), rule=1)  # capi.QUANT_RULE_BATCH_INT
