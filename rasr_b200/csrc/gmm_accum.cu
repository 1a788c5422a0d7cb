// Training E-step on the device: for every aligned frame, the best density of ITS mixture
// only and the accumulation of the frame into that density's estimators -- the scoring work shrinks from M mixtures
// per frame to one, and nothing but the accumulators (read once per pass) leaves the GPU.
//
// Replaces Speech::KeyedEstimator::processAlignedFeature (/root/reference/src/Speech/KeyedEstimator.cc:108-121) ->
//   AssigningScorer::bestDensity(mixture)            GaussDiagonalMaximumFeatureScorer.cc:116-142 (first minimum,
//                                                    f64 sums compared against the f32-stored running best)
//   AbstractMixtureEstimator::accumulate             MixtureEstimator.cc:86-97   (weights_[density] += w)
//   GaussDensityEstimator::accumulate                GaussDensityEstimator.cc:22-32
//   MeanEstimator / CovarianceEstimator accumulators VectorAccumulator.hh:60-67 with std::plus / plusSquare,
//                                                    plusWeighted / plusSquareWeighted (Utilities.hh:110-146), f64
// and, for `viterbi = false`, the posterior-weighted branch of AbstractMixtureSetEstimator::accumulate
// (AbstractMixtureSetEstimator.cc:127-147) with the density posteriors of the sum scorer
// (GaussDiagonalMaximumFeatureScorer.cc:291-298: exp(score - s_k); the maximum scorer has none).
//
// One warp per frame: lane k evaluates density k of the frame's mixture with the reference's operation order (so
// the arg-min is the reference's bit for bit), the warp then replays the reference's sequential scan over the
// K candidates, and lanes d, d+32, .. add x[d] and x[d]^2 into the f64 sums of the chosen mean and covariance.
// The sums are f64 atomics, so their order -- and the last bits of the result -- differ from the reference's
// frame order (tests: 1e-12 relative); counts / weights of unweighted accumulation are exact.  A pooled
// covariance (C = 1) would serialise every frame on one accumulator: each warp keeps its own copy in registers
// and adds it once at the end.
#include <float.h>
#include <math.h>
#include <string.h>

#include "gmm_distance.cuh"
#include "gmm_internal.cuh"

namespace {

constexpr uint32_t kSkip = 0xFFFFFFFFu;

struct AccumParams {
    const float*    x;
    const uint32_t* mixture;  // [T] mixture of every frame, kSkip: not aligned
    const double*   weight;   // [T] or null (1)
    // posterior rows (Baum-Welch alignment): when rowW is set, frame t goes into EVERY mixture emList[emOff[s] + j] of
    // its segment s whose weight rowW[t * ldW + j] is positive, instead of into mixture[t]
    const float*    rowW;
    size_t          ldW;
    const uint32_t *segFrameOff, *emOff, *emList;
    uint32_t        nSeg;
    uint32_t*       best;     // [T] or null: chosen density (index in mixture)
    uint32_t        T, D;
    const float *   mean, *isr, *w2, *logNorm;
    const uint32_t *covIdx, *offsets, *meanIdx;
    double *        meanSum, *meanW, *covSum, *covW, *mixW;
    uint32_t        pooled;
    int             viterbi;
    double          weightThreshold;
};

// SUM = 0: maximum scorer (s_k in f64 against the f32-stored best, GaussDiagonalMaximumFeatureScorer.cc:116-142);
// SUM = 1: sum scorer (s_k = f32(0.5 * (w2 + L + dist)), first minimum, .cc:262-288).
template<int DP, int SUM>
__global__ void __launch_bounds__(256, 4) accumulate_kernel(const AccumParams p) {
    const int      lane = threadIdx.x & 31;
    const uint32_t warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), nWarps = gridDim.x * (blockDim.x >> 5);
    const int      D = (int)p.D, eff = D & ~3;
    double         cov[(DP + 31) / 32], covWeight = 0.0;  // pooled covariance: this warp's share
#pragma unroll
    for (int j = 0; j < (DP + 31) / 32; ++j)
        cov[j] = 0.0;
    // frame t into mixture m with weight w (`weighted`: the weighted accumulation formulas)
    auto process = [&](uint32_t t, uint32_t m, double w, bool weightedItem) {
        const uint32_t a = p.offsets[m], K = p.offsets[m + 1] - a;
        Frames<DP, 1>  f;
#pragma unroll
        for (int d = 0; d < DP; ++d)
            f.x[0][d] = d < D ? p.x[(size_t)t * D + d] : 0.f;
        // lane k of a chunk of 32 densities: w2 + L + dist, as f64 (maximum scorer) and as the sum scorer's f32 s_k
        auto evaluate = [&](uint32_t k0, double& sMax, float& sSum) {
            const uint32_t k = k0 + lane;
            sMax             = 0.0;
            sSum             = 0.f;
            if (k < K) {
                const uint32_t g = a + k, c = p.covIdx[g];
                float          dist;
                distance<DP, 1>(f, reinterpret_cast<const float4*>(p.mean + (size_t)g * DP),
                                reinterpret_cast<const float4*>(p.isr + (size_t)c * DP), D, eff, &dist);
                if (SUM)
                    sSum = (float)(0.5 * (double)__fadd_rn(__fadd_rn(p.w2[g], p.logNorm[c]), dist));
                else
                    sMax = ((double)p.w2[g] + (double)p.logNorm[c]) + (double)dist;
            }
        };
        // frame t into the estimators of density k of the mixture with weight `wt` (all lanes; lane d adds x[d])
        auto accumulate = [&](uint32_t k, double wt, bool weighted) {
            const uint32_t g = a + k, mi = p.meanIdx[g], c = p.covIdx[g];
#pragma unroll
            for (int j = 0; j < (DP + 31) / 32; ++j) {
                const int d = lane + 32 * j;
                if (d < D) {
                    const double xd = (double)p.x[(size_t)t * D + d];
                    // unweighted: sum + x, sum + x*x; weighted: sum + w*x, sum + w*x*x, left to right (Utilities.hh:110-146)
                    atomicAdd(&p.meanSum[(size_t)mi * D + d], weighted ? wt * xd : xd);
                    const double sq = weighted ? wt * xd * xd : xd * xd;
                    if (p.pooled)
                        cov[j] += sq;
                    else
                        atomicAdd(&p.covSum[(size_t)c * D + d], sq);
                }
            }
            if (lane == 0) {
                atomicAdd(&p.mixW[g], wt);
                atomicAdd(&p.meanW[mi], wt);
                if (p.pooled)
                    covWeight += wt;
                else
                    atomicAdd(&p.covW[c], wt);
            }
        };
        // the reference's sequential scan over the densities
        float    bestScore = FLT_MAX;
        uint32_t bestDns   = kSkip;
        for (uint32_t k0 = 0; k0 < K; k0 += 32) {
            double sMax;
            float  sSum;
            evaluate(k0, sMax, sSum);
            const uint32_t n = min(32u, K - k0);
            for (uint32_t j = 0; j < n; ++j) {
                if (SUM) {
                    const float sj = __shfl_sync(0xFFFFFFFFu, sSum, j);
                    if (bestScore > sj) {
                        bestScore = sj;
                        bestDns   = k0 + j;
                    }
                }
                else {
                    const double sj = __shfl_sync(0xFFFFFFFFu, sMax, j);
                    if ((double)bestScore > sj) {
                        bestScore = (float)sj;
                        bestDns   = k0 + j;
                    }
                }
            }
        }
        if (p.best && lane == 0)
            p.best[t] = bestDns;
        if (p.viterbi) {
            if (bestDns != kSkip)  // no density below FLT_MAX: nothing the reference could index either
                accumulate(bestDns, w, weightedItem);
        }
        else if (SUM) {
            // score = best - log sum_k exp(best - s_k), posteriors exp(score - s_k) (.cc:276-298), each density
            // with weight w * posterior above the threshold (AbstractMixtureSetEstimator.cc:133-146)
            float sumExp = 0.f;
            for (uint32_t k0 = 0; k0 < K; k0 += 32) {
                double sMax;
                float  sSum;
                evaluate(k0, sMax, sSum);
                const uint32_t n = min(32u, K - k0);
                for (uint32_t j = 0; j < n; ++j)
                    sumExp += expf(bestScore - __shfl_sync(0xFFFFFFFFu, sSum, j));
            }
            const float score = bestScore - logf(sumExp);
            for (uint32_t k0 = 0; k0 < K; k0 += 32) {
                double sMax;
                float  sSum;
                evaluate(k0, sMax, sSum);
                const uint32_t n = min(32u, K - k0);
                for (uint32_t j = 0; j < n; ++j) {
                    const double fin = w * (double)expf(score - __shfl_sync(0xFFFFFFFFu, sSum, j));
                    if (fin > p.weightThreshold)
                        accumulate(k0 + j, fin, true);
                }
            }
        }
    };
    if (p.rowW) {
        const float itemCut = (float)fmax(0.0, p.weightThreshold);
        for (uint32_t t = warp; t < p.T; t += nWarps) {
            uint32_t lo = 0, hi = p.nSeg;  // segment of frame t: last s with segFrameOff[s] <= t
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (p.segFrameOff[mid] <= t)
                    lo = mid;
                else
                    hi = mid;
            }
            const uint32_t e0 = p.emOff[lo], n = p.emOff[lo + 1] - e0;
            const float*   row = p.rowW + (size_t)t * p.ldW;
            for (uint32_t j0 = 0; j0 < n; j0 += 32) {
                const float wj   = j0 + lane < n ? row[j0 + lane] : 0.f;
                // items at or below the estimators' weight threshold are dropped (the reference's knob for this is the
                // aligner's min-weight, default 0: create the accumulators with threshold 0 for the literal behaviour)
                uint32_t    mask = __ballot_sync(0xFFFFFFFFu, wj > itemCut);
                while (mask) {
                    const int j = __ffs(mask) - 1;
                    mask &= mask - 1;
                    process(t, p.emList[e0 + j0 + j], (double)__shfl_sync(0xFFFFFFFFu, wj, j), true);
                }
            }
        }
    }
    else {
        for (uint32_t t = warp; t < p.T; t += nWarps) {
            const uint32_t m = p.mixture[t];
            if (m != kSkip)
                process(t, m, p.weight ? p.weight[t] : 1.0, p.weight != nullptr);
        }
    }
    if (p.pooled) {
#pragma unroll
        for (int j = 0; j < (DP + 31) / 32; ++j) {
            const int d = lane + 32 * j;
            if (d < D && cov[j] != 0.0)
                atomicAdd(&p.covSum[d], cov[j]);
        }
        if (lane == 0 && covWeight != 0.0)
            atomicAdd(&p.covW[0], covWeight);
    }
}

template<int DP>
int launch(gmm_handle* h, const AccumParams& p) {
    const int grid = h->smCount * 8;
    h->timeBegin();
    if (h->mode == GMM_MODE_SUM)
        accumulate_kernel<DP, 1><<<grid, 256, 0, h->stream>>>(p);
    else
        accumulate_kernel<DP, 0><<<grid, 256, 0, h->stream>>>(p);
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    return GMM_OK;
}

}  // namespace

static int accumRun(gmm_handle* h, AccumParams& p) {
    const DirectModel& dm = h->direct;
    const AccumModel&  am = h->accum;
    p.D               = h->D;
    p.mean            = dm.mean;
    p.isr             = dm.isr;
    p.w2              = dm.w2;
    p.logNorm         = dm.logNorm;
    p.covIdx          = dm.covIdx;
    p.offsets         = dm.offsets;
    p.meanIdx         = am.meanIdx;
    p.meanSum         = am.meanSum;
    p.meanW           = am.meanW;
    p.covSum          = am.covSum;
    p.covW            = am.covW;
    p.mixW            = am.mixW;
    p.pooled          = h->C == 1;
    p.viterbi         = am.viterbi;
    p.weightThreshold = am.weightThreshold;
    switch (dm.DP) {
    case 16: return launch<16>(h, p);
    case 32: return launch<32>(h, p);
    case 48: return launch<48>(h, p);
    case 64: return launch<64>(h, p);
    case 96: return launch<96>(h, p);
    case 128: return launch<128>(h, p);
    default:
        h->setError("accumulate kernel: feature dimension above 128 is not supported");
        return GMM_ERR_UNSUPPORTED;
    }
}

int gmm_accum_launch(gmm_handle* h, uint32_t T, const float* d_x, const uint32_t* d_mixture, const double* d_weight,
                     uint32_t* d_best) {
    if (T == 0)
        return GMM_OK;
    AccumParams p;
    memset(&p, 0, sizeof(p));
    p.x       = d_x;
    p.mixture = d_mixture;
    p.weight  = d_weight;
    p.best    = d_best;
    p.T       = T;
    return accumRun(h, p);
}

int gmm_accum_rows_launch(gmm_handle* h, uint32_t nSeg, uint32_t T, const float* d_x, const uint32_t* d_frameOff,
                          const uint32_t* d_emOff, const uint32_t* d_emList, const float* d_rowWeights, size_t ldW) {
    if (T == 0 || nSeg == 0)
        return GMM_OK;
    AccumParams p;
    memset(&p, 0, sizeof(p));
    p.x           = d_x;
    p.T           = T;
    p.rowW        = d_rowWeights;
    p.ldW         = ldW;
    p.segFrameOff = d_frameOff;
    p.emOff       = d_emOff;
    p.emList      = d_emList;
    p.nSeg        = nSeg;
    return accumRun(h, p);
}
