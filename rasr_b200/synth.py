"""Synthetic mixture sets and feature streams of the shapes BASELINE.json names.

numpy ``default_rng`` with fixed seeds so the CPU oracle, the compiled reference and the GPU path read
identical bytes.  No scoring arithmetic lives here.
"""
from __future__ import annotations

import numpy as np

from .gms import MixtureSetData

F64_MIN = -np.finfo(np.float64).max  # Core::Type<f64>::min, the log-weight of a zero-weight density


def make_model(M: int, K, D: int, *, seed: int = 1001, covariance: str = "pooled", ragged: bool = False,
               shared: bool = False, zero_weight_frac: float = 0.0, mean_scale: float = 2.0) -> MixtureSetData:
    """Random diagonal-covariance mixture set.

    covariance: "pooled" (C=1, the RASR default, Mm/Module.cc:59-60), "density" (one per density) or
    "mixture" (one per mixture).  ragged: densities per mixture uniform in [1, K].  shared: some
    mixtures reference densities of other mixtures (tying, MixtureSet.hh:47-50).
    """
    rng = np.random.default_rng(seed)
    if ragged:
        ks = rng.integers(1, K + 1, size=M)
    else:
        ks = np.full(M, K, dtype=np.int64)
    offs = np.zeros(M + 1, dtype=np.uint32)
    offs[1:] = np.cumsum(ks)
    G = int(offs[-1])
    n_dens = G
    mix_density = np.arange(G, dtype=np.uint32)
    if shared and G > 4:
        # re-point ~10 % of the slots at an earlier density: exercises tying without changing G
        pick = rng.random(G) < 0.1
        mix_density[pick] = rng.integers(0, n_dens, size=int(pick.sum()), dtype=np.uint32)
    if covariance == "pooled":
        n_cov = 1
        dens_cov = np.zeros(n_dens, dtype=np.uint32)
    elif covariance == "density":
        n_cov = n_dens
        dens_cov = np.arange(n_dens, dtype=np.uint32)
    elif covariance == "mixture":
        n_cov = M
        dens_cov = np.repeat(np.arange(M, dtype=np.uint32), ks)
    else:
        raise ValueError(covariance)
    variances = np.exp(rng.uniform(np.log(0.25), np.log(4.0), size=(n_cov, D))).astype(np.float32)
    sigma = np.sqrt(variances.astype(np.float64))
    means = (rng.standard_normal((n_dens, D)) * sigma[dens_cov] * mean_scale).astype(np.float32)
    dens_mean = np.arange(n_dens, dtype=np.uint32)
    # Dirichlet(1_K) weights per mixture, stored as f64 log-weights
    e = rng.exponential(size=G)
    logw = np.empty(G, dtype=np.float64)
    for m in range(M):
        a, b = int(offs[m]), int(offs[m + 1])
        logw[a:b] = np.log(e[a:b] / e[a:b].sum())
    if zero_weight_frac > 0:
        z = rng.random(G) < zero_weight_frac
        # keep at least the first density of each mixture alive
        z[offs[:-1][ks > 0]] = False
        logw[z] = F64_MIN
    return MixtureSetData(D, offs, mix_density, logw, dens_mean, dens_cov, means, variances)


def make_features(model: MixtureSetData, T: int, *, seed: int = 2001) -> np.ndarray:
    """x_t = mu_{g_t} + sigma (.) N(0,1) with g_t uniform over densities: frames lie near some density,
    so scores are realistic (about D/2 .. a few hundred) and arg-mins are non-trivial."""
    rng = np.random.default_rng(seed)
    g = rng.integers(0, model.n_densities, size=T)
    sigma = np.sqrt(model.variances.astype(np.float64))[model.dens_cov[g]]
    x = model.means[model.dens_mean[g]].astype(np.float64) + sigma * rng.standard_normal((T, model.D))
    return np.ascontiguousarray(x, dtype=np.float32)


# The configurations of BASELINE.json, by name.
CONFIGS = {
    "cfg1": dict(M=4501, K=16, D=45, T=1000, covariance="pooled", seed=1001),
    "cfg2": dict(M=4501, K=16, D=45, T=1000, covariance="pooled", seed=1001),
    "cfg3": dict(M=12001, K=64, D=45, T=10000, covariance="pooled", seed=1001),
    "cfg5": dict(M=65536, K=64, D=45, T=1024, covariance="pooled", seed=1001),
}


def make_config(name: str, T: int | None = None):
    c = dict(CONFIGS[name])
    t = c.pop("T")
    model = make_model(c["M"], c["K"], c["D"], seed=c["seed"], covariance=c["covariance"])
    return model, make_features(model, T if T is not None else t)
