"""Generates tests/golden/*.npz by running the UNMODIFIED reference scorer classes (oracle/_ref/ref_scorer,
built from /root/reference by oracle/ref_build/Makefile) on small synthetic models.  Run in the build
container only:  python tests/golden/make_golden.py

Each fixture holds the raw model tables, the features, and the reference's outputs:
  max_score/max_best   Mm::GaussDiagonalMaximumFeatureScorer     score(m), bestDensity(m)
  sum_score/sum_best   Mm::GaussDiagonalSumFeatureScorer
  simd_*               Mm::SimdGaussDiagonalMaximumFeatureScorer: final f32 score, q (s32), best, and the
                       host tables its init() produced (qmean, qconst, qcov, isrs, scalingSquared) + qx
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from rasr_b200 import synth  # noqa: E402

CASES = {
    # name: (make_model kwargs, T)
    "pooled_k16_d45": (dict(M=7, K=16, D=45, seed=11), 9),
    "pooled_k1_d1": (dict(M=5, K=1, D=1, seed=12), 6),
    "pooled_k3_d4": (dict(M=6, K=3, D=4, seed=13), 7),
    "pooled_k5_d15": (dict(M=4, K=5, D=15, seed=14), 5),
    "pooled_k8_d16": (dict(M=4, K=8, D=16, seed=15), 5),
    "pooled_k8_d17": (dict(M=4, K=8, D=17, seed=16), 5),
    "density_cov_ragged_shared": (dict(M=9, K=6, D=13, seed=17, covariance="density", ragged=True, shared=True), 8),
    "mixture_cov_zero_weights": (dict(M=8, K=7, D=21, seed=18, covariance="mixture", zero_weight_frac=0.3), 8),
    "pooled_far_means": (dict(M=5, K=4, D=10, seed=19, mean_scale=40.0), 6),
}


def main():
    assert O.build_ref() and O.ref_available(), "needs /root/reference (build container)"
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, (kw, T) in CASES.items():
        ms = synth.make_model(**kw)
        x = synth.make_features(ms, T, seed=100 + kw["seed"])
        if name == "pooled_far_means":
            x[0] *= 1e3  # saturates the u8 quantizer
        r_max = O.run_ref(ms, x, "max")
        r_sum = O.run_ref(ms, x, "sum")
        r_simd = O.run_ref(ms, x, "simd")
        np.savez_compressed(
            os.path.join(out_dir, name + ".npz"), D=ms.D, mix_offsets=ms.mix_offsets, mix_density=ms.mix_density,
            mix_logw=ms.mix_logw, dens_mean=ms.dens_mean, dens_cov=ms.dens_cov, means=ms.means,
            variances=ms.variances, x=x, max_score=r_max["score"], max_best=r_max["best"], sum_score=r_sum["score"],
            sum_best=r_sum["best"], simd_score=r_simd["score"], simd_q=r_simd["q"], simd_best=r_simd["best"],
            simd_qmean=r_simd["qmean"], simd_qconst=r_simd["qconst"], simd_qcov=r_simd["qcov"],
            simd_isrs=r_simd["isrs"], simd_qx=r_simd["qx"], simd_scaling_squared=r_simd["scaling_squared"])
        print(name, "ok", r_max["score"].shape)


if __name__ == "__main__":
    main()
