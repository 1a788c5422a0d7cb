"""Generates tests/golden/est_*.npz by running the UNMODIFIED reference text writer / reader and estimator stack
(oracle/_ref/ref_estimator, built from /root/reference by oracle/ref_build/Makefile: Mm/MixtureSet.cc:142-214,
Mm/AbstractMixtureSetEstimator.cc:117-147,299-335,433-506, MixtureSetEstimator.cc, MixtureEstimator.cc,
GaussDensityEstimator.cc) on small synthetic models.  Run in the build container only:
    python tests/golden/make_estimator_golden.py

Each fixture holds the raw model tables, features, the aligned mixture and weight of every frame, and the reference's
files as byte arrays:
  pms                      MixtureSet::write of the model (text, "#Version: 2.0")
  mixset_<case>, est_<case>, best_<case>   the binary accumulator file ("MIXSET"), estimate() of it as text, and (viterbi)
                           the density every frame was assigned to, for the cases
     vit      viterbi, maximum scorer, unweighted accumulate(mixture, feature)
     vitw     viterbi, maximum scorer, accumulate(mixture, feature, weight)
     vitsum   viterbi, sum scorer, weighted
     bw       baum-welch (density posteriors of the sum scorer), weighted, weight threshold f32 epsilon
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from rasr_b200 import synth  # noqa: E402

MODELS = {
    "est_pooled_k16_d45": (dict(M=6, K=16, D=45, seed=21), 700),
    "est_mixture_cov_shared": (dict(M=5, K=3, D=7, seed=22, covariance="mixture", shared=True), 300),
    "est_density_cov_ragged": (dict(M=7, K=5, D=13, seed=23, covariance="density", ragged=True), 400),
}
CASES = {"vit": ("viterbi", "max", False), "vitw": ("viterbi", "max", True), "vitsum": ("viterbi", "sum", True),
         "bw": ("baum-welch", "sum", True)}


def main():
    assert O.build_ref() and O.ref_estimator_available(), "needs /root/reference (build container)"
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, (kw, T) in MODELS.items():
        ms = synth.make_model(**kw)
        x = synth.make_features(ms, T, seed=200 + kw["seed"])
        rng = np.random.default_rng(300 + kw["seed"])
        # runs of frames on one mixture, as alignments have them; a few frames skipped
        mixture = np.repeat(rng.integers(0, ms.M, T // 4 + 1), 4)[:T].astype(np.uint32)
        mixture[: ms.M * 8] = np.repeat(np.arange(ms.M), 8)   # every mixture is observed
        mixture[rng.integers(0, T, 5)] = 0xFFFFFFFF
        weight = rng.uniform(0.05, 2.0, T)
        fix = dict(D=ms.D, mix_offsets=ms.mix_offsets, mix_density=ms.mix_density, mix_logw=ms.mix_logw,
                   dens_mean=ms.dens_mean, dens_cov=ms.dens_cov, means=ms.means, variances=ms.variances, x=x,
                   mixture=mixture, weight=weight)
        for case, (mode, scorer, weighted) in CASES.items():
            r = O.run_ref_estimator(ms, x, mixture, weight if weighted else None, mode=mode, scorer=scorer)
            assert r["pms"] == r["reread_pms"] and r["mixset"] == r["mixset2"], "the reference does not round-trip"
            fix["pms"] = np.frombuffer(r["pms"], np.uint8)
            fix["mixset_" + case] = np.frombuffer(r["mixset"], np.uint8)
            fix["est_" + case] = np.frombuffer(r["est_pms"], np.uint8)
            fix["best_" + case] = r["best"]
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **fix)
        print(name, "ok", {k: v.shape for k, v in fix.items() if k.startswith("mixset")})


if __name__ == "__main__":
    main()
