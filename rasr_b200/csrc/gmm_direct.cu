// Direct ("reference-order") float kernel: one lane = one frame (two with FR=2), its feature vector held in
// registers; a warp walks the densities of a mixture in order, every lane reading the same mean / inverse
// std-dev words (uniform, L1-resident 128-bit loads).  The distance is evaluated with the operation order of
// the reference's SSE3 code -- four lane sums over d < 4*floor(D/4), (l0+l1)+(l2+l3), then the scalar tail,
// no FMA contraction (/root/reference/src/Mm/GaussDiagonalMaximumFeatureScorer.cc:144-218) -- and the
// mixture minimum with its f64 adds and f32-stored running best (.cc:116-142), so the maximum-approximation
// scores are bit-identical to the reference's whenever the tables are.  It is the engine for models the
// tensor kernel does not take (K > 128 per mixture, D > 128 ...) and the cross-check of the tensor kernel.
#include <float.h>
#include <math.h>

#include "gmm_distance.cuh"
#include "gmm_internal.cuh"

namespace {

// MODE 0 = max, 1 = sum, 2 = posteriors of the sum scorer (reads scores, writes post[T x G])
template<int DP, int FR, int MODE>
__global__ void __launch_bounds__(256) direct_kernel(const float* __restrict__ x, int T, int D, int M,
                                                     const float* __restrict__ mean, const float* __restrict__ isr,
                                                     const uint32_t* __restrict__ covIdx,
                                                     const float* __restrict__ w2, const float* __restrict__ logNorm,
                                                     const uint32_t* __restrict__ offsets, int mixPerBlock,
                                                     float* __restrict__ scores, uint32_t* __restrict__ best,
                                                     size_t ld, float* __restrict__ post, size_t G) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t0   = (blockIdx.y * 8 + warp) * 32 * FR + lane;
    if ((blockIdx.y * 8 + warp) * 32 * FR >= T)
        return;  // whole warp out of range
    Frames<DP, FR> f;
#pragma unroll
    for (int r = 0; r < FR; ++r) {
        const int t = t0 + 32 * r;
#pragma unroll
        for (int d = 0; d < DP; ++d)
            f.x[r][d] = (t < T && d < D) ? x[(size_t)t * D + d] : 0.f;
    }
    const int eff  = D & ~3;
    const int mBeg = blockIdx.x * mixPerBlock;
    const int mEnd = min(M, mBeg + mixPerBlock);
    for (int m = mBeg; m < mEnd; ++m) {
        const uint32_t a = offsets[m], b = offsets[m + 1];
        if (MODE == 0) {
            float    bestScore[FR];
            uint32_t bestDns[FR];
#pragma unroll
            for (int r = 0; r < FR; ++r) {
                bestScore[r] = FLT_MAX;
                bestDns[r]   = 0xFFFFFFFFu;
            }
            for (uint32_t g = a; g < b; ++g) {
                const uint32_t c = covIdx[g];
                float          dist[FR];
                distance<DP, FR>(f, reinterpret_cast<const float4*>(mean + (size_t)g * DP),
                                 reinterpret_cast<const float4*>(isr + (size_t)c * DP), D, eff, dist);
                const double base = (double)w2[g] + (double)logNorm[c];
#pragma unroll
                for (int r = 0; r < FR; ++r) {
                    const double s = base + (double)dist[r];
                    if ((double)bestScore[r] > s) {
                        bestScore[r] = (float)s;
                        bestDns[r]   = g - a;
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < FR; ++r) {
                const int t = t0 + 32 * r;
                if (t < T) {
                    scores[(size_t)t * ld + m] = (float)(0.5 * (double)bestScore[r]);
                    if (best)
                        best[(size_t)t * ld + m] = bestDns[r];
                }
            }
        }
        else {
            // s_k = f32(0.5 * (w2 + L + dist)); best = first minimum; score = best - log sum exp(best - s_k)
            float    runBest[FR], sumExp[FR], fin[FR];
            uint32_t bestDns[FR];
#pragma unroll
            for (int r = 0; r < FR; ++r) {
                runBest[r] = FLT_MAX;
                sumExp[r]  = 0.f;
                bestDns[r] = 0xFFFFFFFFu;
                const int t = t0 + 32 * r;
                fin[r]      = (MODE == 2 && t < T) ? scores[(size_t)t * ld + m] : 0.f;
            }
            for (uint32_t g = a; g < b; ++g) {
                const uint32_t c = covIdx[g];
                float          dist[FR];
                distance<DP, FR>(f, reinterpret_cast<const float4*>(mean + (size_t)g * DP),
                                 reinterpret_cast<const float4*>(isr + (size_t)c * DP), D, eff, dist);
                const float base = __fadd_rn(w2[g], logNorm[c]);
#pragma unroll
                for (int r = 0; r < FR; ++r) {
                    const float s = (float)(0.5 * (double)__fadd_rn(base, dist[r]));
                    if (MODE == 2) {
                        const int t = t0 + 32 * r;
                        if (t < T)
                            post[(size_t)t * G + g] = expf(fin[r] - s);
                    }
                    else if (runBest[r] > s) {
                        // online rescale: sum exp(best - s_k) with the running best
                        sumExp[r]  = sumExp[r] * expf(s - runBest[r]) + 1.f;
                        runBest[r] = s;
                        bestDns[r] = g - a;
                    }
                    else {
                        sumExp[r] += expf(runBest[r] - s);
                    }
                }
            }
            if (MODE == 1) {
#pragma unroll
                for (int r = 0; r < FR; ++r) {
                    const int t = t0 + 32 * r;
                    if (t < T) {
                        scores[(size_t)t * ld + m] = runBest[r] - logf(sumExp[r]);
                        if (best)
                            best[(size_t)t * ld + m] = bestDns[r];
                    }
                }
            }
        }
    }
}

template<int DP, int FR>
int launch(gmm_handle* h, int mode, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld,
           float* d_post) {
    const DirectModel& dm = h->direct;
    const int framesPerBlock = 8 * 32 * FR;
    const int gy             = (T + framesPerBlock - 1) / framesPerBlock;
    int       wantX          = (4 * h->smCount + gy - 1) / gy;
    if (wantX < 1)
        wantX = 1;
    int mixPerBlock = (h->M + wantX - 1) / wantX;
    if (mixPerBlock < 1)
        mixPerBlock = 1;
    const int gx = (h->M + mixPerBlock - 1) / mixPerBlock;
    dim3      grid(gx, gy), block(256);
#define GMM_ARGS d_x, (int)T, (int)h->D, (int)h->M, dm.mean, dm.isr, dm.covIdx, dm.w2, dm.logNorm, \
    dm.offsets, mixPerBlock, d_scores, d_best, ld, d_post, (size_t)h->G
    h->timeBegin();
    if (mode == 0)
        direct_kernel<DP, FR, 0><<<grid, block, 0, h->stream>>>(GMM_ARGS);
    else if (mode == 1)
        direct_kernel<DP, FR, 1><<<grid, block, 0, h->stream>>>(GMM_ARGS);
    else
        direct_kernel<DP, FR, 2><<<grid, block, 0, h->stream>>>(GMM_ARGS);
#undef GMM_ARGS
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    return GMM_OK;
}

int dispatch(gmm_handle* h, int mode, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld,
             float* d_post) {
    if (T == 0)
        return GMM_OK;
    switch (h->direct.DP) {
    case 16: return launch<16, 2>(h, mode, T, d_x, d_scores, d_best, ld, d_post);
    case 32: return launch<32, 2>(h, mode, T, d_x, d_scores, d_best, ld, d_post);
    case 48: return launch<48, 2>(h, mode, T, d_x, d_scores, d_best, ld, d_post);
    case 64: return launch<64, 2>(h, mode, T, d_x, d_scores, d_best, ld, d_post);
    case 96: return launch<96, 1>(h, mode, T, d_x, d_scores, d_best, ld, d_post);
    case 128: return launch<128, 1>(h, mode, T, d_x, d_scores, d_best, ld, d_post);
    default:
        h->setError("direct kernel: feature dimension above 128 is not supported");
        return GMM_ERR_UNSUPPORTED;
    }
}

// Density posteriors of ONE mixture per frame (the training consumer asks for the mixture a frame is aligned to):
// GaussDiagonalSumFeatureScorer::calculateScoresAndNumberOfDensities / calculateScoreAndDensity /
// calculateDensityPosteriorProbabilities (GaussDiagonalMaximumFeatureScorer.cc:239-298) in the reference's order:
// s_k = 0.5f * (w2 + L + dist) in f32 with the 4-lane distance, best = first minimum, sumExp accumulated density by
// density, score = best - log(sumExp), p_k = exp(score - s_k).  One CTA per frame, one thread per density for the
// distances, thread 0 for the two sequential passes (K is small; this is not a throughput path).
// post [T x ldP]: columns >= the mixture's density count are left untouched; frames with mixture 0xFFFFFFFF are skipped.
__global__ void __launch_bounds__(128) mixture_posteriors_kernel(const float* __restrict__ x, int T, int D, int DP,
                                                                 const float* __restrict__ mean, const float* __restrict__ isr,
                                                                 const uint32_t* __restrict__ covIdx,
                                                                 const float* __restrict__ w2,
                                                                 const float* __restrict__ logNorm,
                                                                 const uint32_t* __restrict__ offsets,
                                                                 const uint32_t* __restrict__ mixture, int M,
                                                                 float* __restrict__ post, size_t ldP,
                                                                 float* __restrict__ score) {
    extern __shared__ float sc[];  // density scores of the frame's mixture
    const int t = blockIdx.x;
    if (t >= T)
        return;
    const uint32_t m = mixture[t];
    if (m >= (uint32_t)M)
        return;
    const uint32_t a = offsets[m], b = offsets[m + 1];
    const float*   xt  = x + (size_t)t * D;
    const int      eff = D & ~3;
    for (uint32_t g = a + threadIdx.x; g < b; g += blockDim.x) {
        const uint32_t c  = covIdx[g];
        const float*   mu = mean + (size_t)g * DP;
        const float*   is = isr + (size_t)c * DP;
        float          l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
        for (int d = 0; d < eff; d += 4) {
            const float d0 = __fmul_rn(__fsub_rn(mu[d], xt[d]), is[d]);
            const float d1 = __fmul_rn(__fsub_rn(mu[d + 1], xt[d + 1]), is[d + 1]);
            const float d2 = __fmul_rn(__fsub_rn(mu[d + 2], xt[d + 2]), is[d + 2]);
            const float d3 = __fmul_rn(__fsub_rn(mu[d + 3], xt[d + 3]), is[d + 3]);
            l0 = __fadd_rn(l0, __fmul_rn(d0, d0));
            l1 = __fadd_rn(l1, __fmul_rn(d1, d1));
            l2 = __fadd_rn(l2, __fmul_rn(d2, d2));
            l3 = __fadd_rn(l3, __fmul_rn(d3, d3));
        }
        float dist = __fadd_rn(0.f, __fadd_rn(__fadd_rn(l0, l1), __fadd_rn(l2, l3)));
        for (int d = eff; d < D; ++d) {
            const float dd = __fmul_rn(__fsub_rn(mu[d], xt[d]), is[d]);
            dist           = __fadd_rn(dist, __fmul_rn(dd, dd));
        }
        const float s = __fadd_rn(__fadd_rn(w2[g], logNorm[c]), dist);
        sc[g - a]     = __fmul_rn(0.5f, s);
    }
    __syncthreads();
    __shared__ float fin;
    if (threadIdx.x == 0) {
        float best = 3.402823466e+38f;
        for (uint32_t k = 0; k < b - a; ++k)
            if (best > sc[k])
                best = sc[k];
        float sumExp = 0.f;
        for (uint32_t k = 0; k < b - a; ++k)
            sumExp += expf(best - sc[k]);
        fin = best - logf(sumExp);
        if (score)
            score[t] = fin;
    }
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < b - a; k += blockDim.x)
        post[(size_t)t * ldP + k] = expf(fin - sc[k]);
}

}  // namespace

int gmm_direct_mixture_posteriors(gmm_handle* h, uint32_t T, const float* d_x, const uint32_t* d_mixture, float* d_post,
                                  size_t ldP, float* d_score, uint32_t maxK) {
    if (T == 0)
        return GMM_OK;
    const DirectModel& dm = h->direct;
    mixture_posteriors_kernel<<<T, 128, (size_t)std::max(1u, maxK) * sizeof(float), h->stream>>>(
            d_x, (int)T, (int)h->D, (int)dm.DP, dm.mean, dm.isr, dm.covIdx, dm.w2, dm.logNorm, dm.offsets, d_mixture,
            (int)h->M, d_post, ldP, d_score);
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    return GMM_OK;
}

int gmm_direct_score(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld) {
    return dispatch(h, h->mode == GMM_MODE_SUM ? 1 : 0, T, d_x, d_scores, d_best, ld, nullptr);
}

int gmm_direct_posteriors(gmm_handle* h, uint32_t T, const float* d_x, float* d_post) {
    // pass 1: scores of the sum scorer into scratch, pass 2: exp(score - s_k)
    void* sc = nullptr;
    int   rc = h->reserve(SCR_SCORE, (size_t)T * h->M * sizeof(float), &sc);
    if (rc)
        return rc;
    rc = dispatch(h, 1, T, d_x, (float*)sc, nullptr, h->M, nullptr);
    if (rc)
        return rc;
    return dispatch(h, 2, T, d_x, (float*)sc, nullptr, h->M, d_post);
}
