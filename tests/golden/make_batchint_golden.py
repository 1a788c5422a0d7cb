"""Generates tests/golden/batchint_*.npz: the UNMODIFIED reference's Mm::BatchIntFeatureScorer
(`batch-diagonal-maximum-int`, src/Mm/BatchFeatureScorer.cc:300-472) run by oracle/_ref/ref_scorer through its
buffered protocol, on the pooled-covariance models of make_golden.py plus one with zero weights (whose constants hit
the 0x80000000 conversion).  Run in the build container only:  python tests/golden/make_batchint_golden.py

Each fixture: the raw model tables, the features x, and `score` [T x M] f32 as the reference answered it."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from rasr_b200 import synth  # noqa: E402

CASES = {
    "batchint_pooled_k16_d45": (dict(M=7, K=16, D=45, seed=11), 9),
    "batchint_pooled_k1_d1": (dict(M=5, K=1, D=1, seed=12), 6),
    "batchint_pooled_k8_d17": (dict(M=4, K=8, D=17, seed=16), 5),
    "batchint_pooled_far_means": (dict(M=5, K=4, D=10, seed=19, mean_scale=40.0), 6),
    "batchint_pooled_ragged_zero_weights": (dict(M=40, K=9, D=33, seed=5, zero_weight_frac=0.2, ragged=True), 50),
}


def main():
    assert O.build_ref() and O.ref_available(), "needs /root/reference (build container)"
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, (kw, T) in CASES.items():
        ms = synth.make_model(**kw)
        x = synth.make_features(ms, T, seed=100 + kw["seed"])
        if name.endswith("far_means"):
            x[0] *= 1e3  # saturates the u8 quantizer
        r = O.run_ref(ms, x, "batchint")
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), D=ms.D, mix_offsets=ms.mix_offsets,
                            mix_density=ms.mix_density, mix_logw=ms.mix_logw, dens_mean=ms.dens_mean, dens_cov=ms.dens_cov,
                            means=ms.means, variances=ms.variances, x=x, score=r["score"])
        print(name, "ok", r["score"].shape)


if __name__ == "__main__":
    main()
