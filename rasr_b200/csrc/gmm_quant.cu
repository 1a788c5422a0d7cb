// Quantized (u8) scorer, bit-exact with the reference's SIMD-diagonal-maximum scorer.
//
//   qx[t][c][d] = Q(x[t][d] * isrS[c][d]),  Q(v) = clip((int)round(v) + 128, 0, 255), zero padded to Dpad
//                 (/root/reference/src/Mm/SimdFeatureScorer.cc:22-35, IntelOptimization.cc:50-67, Utilities.hh:179-191)
//   q[t][m]     = min_k( qconst[k] + sum_{d<Dpad} (qmean[k][d] - qx[t][cov(k)][d])^2 ), first strict minimum
//                 (SimdFeatureScorer.cc:158-176; the JIT SSE2 loop SSE2CodeGenerator.cc:324-374)
//   score[t][m] = f32( 0.5 * q / scalingSquared )                      (SimdFeatureScorer.cc:135-145)
//
// The squared distance is evaluated as |a|^2 + |b|^2 - 2 a.b with dp4a (u8 x u8 -> u32): every partial is an
// exact integer below 2^22, and all adds wrap modulo 2^32 exactly like the reference's int arithmetic, so
// the s32 results are identical -- including the reference's behaviour for a zero-weight density, whose
// constant is the x86 "integer indefinite" 0x80000000.
//
// One lane = FR frames with their quantized vectors in registers; the warp walks the densities of a mixture
// in order reading each quantized mean once (uniform 128-bit loads) for 32*FR frames.
#include "gmm_internal.cuh"

#include <algorithm>
#include <mutex>

namespace {

__device__ __forceinline__ uint8_t quantize_u8(float v) {
    // round(): half away from zero; (int) of an out-of-range value is cvttss2si's 0x80000000
    float r = roundf(v);
    int   i = (fabsf(r) < 2147483648.f) ? __float2int_rz(r) : (int)0x80000000;
    int   s = (int)((unsigned)i + 128u);
    s       = s < 0 ? 0 : (s > 255 ? 255 : s);
    return (uint8_t)s;
}

__global__ void quantize_kernel(const float* __restrict__ x, int T, int D, int Dpad, int C,
                                const float* __restrict__ isrS, uint8_t* __restrict__ qx) {
    // one thread per (t, c, 4 consecutive dims)
    const int    W     = Dpad / 4;
    const size_t total = (size_t)T * C * W;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int    w  = (int)(i % W);
        const size_t tc = i / W;
        const int    c  = (int)(tc % C);
        const size_t t  = tc / C;
        uint32_t     pk = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int d = 4 * w + j;
            uint8_t   v = 0;
            if (d < D)
                v = quantize_u8(__fmul_rn(x[t * D + d], isrS[(size_t)c * D + d]));
            pk |= (uint32_t)v << (8 * j);
        }
        reinterpret_cast<uint32_t*>(qx)[i] = pk;
    }
}

template<int DW, int FR, bool POOLED>
__global__ void __launch_bounds__(256) quant_score_kernel(const uint8_t* __restrict__ qx, int T, int M, int C,
                                                          const uint8_t* __restrict__ qmean,
                                                          const uint32_t* __restrict__ covIdx,
                                                          const int32_t* __restrict__ qconst2,
                                                          const uint32_t* __restrict__ offsets, int mixPerBlock,
                                                          double scalingSquared, int32_t* __restrict__ q,
                                                          uint32_t* __restrict__ best, float* __restrict__ score,
                                                          size_t ld) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tw   = (blockIdx.y * 8 + warp) * 32 * FR;
    if (tw >= T)
        return;
    const int t0 = tw + lane;
    uint32_t  fx[FR][DW];
    uint32_t  xx[FR];
    if (POOLED) {
#pragma unroll
        for (int r = 0; r < FR; ++r) {
            const int t = min(t0 + 32 * r, T - 1);
            xx[r]       = 0;
#pragma unroll
            for (int w4 = 0; w4 < DW / 4; ++w4) {
                uint4 v = reinterpret_cast<const uint4*>(qx + (size_t)t * DW * 4)[w4];
                fx[r][4 * w4 + 0] = v.x;
                fx[r][4 * w4 + 1] = v.y;
                fx[r][4 * w4 + 2] = v.z;
                fx[r][4 * w4 + 3] = v.w;
            }
#pragma unroll
            for (int w = 0; w < DW; ++w)
                xx[r] = __dp4a(fx[r][w], fx[r][w], xx[r]);
        }
    }
    const int mBeg = blockIdx.x * mixPerBlock;
    const int mEnd = min(M, mBeg + mixPerBlock);
    for (int m = mBeg; m < mEnd; ++m) {
        const uint32_t a = offsets[m], b = offsets[m + 1];
        int32_t        minScore[FR];
        uint32_t       bestDns[FR];
#pragma unroll
        for (int r = 0; r < FR; ++r) {
            minScore[r] = 0x7FFFFFFF;
            bestDns[r]  = 0xFFFFFFFFu;
        }
        for (uint32_t g = a; g < b; ++g) {
            uint32_t mw[DW];
#pragma unroll
            for (int w4 = 0; w4 < DW / 4; ++w4) {
                uint4 v = __ldg(reinterpret_cast<const uint4*>(qmean + (size_t)g * DW * 4) + w4);
                mw[4 * w4 + 0] = v.x;
                mw[4 * w4 + 1] = v.y;
                mw[4 * w4 + 2] = v.z;
                mw[4 * w4 + 3] = v.w;
            }
            const uint32_t c2 = (uint32_t)qconst2[g];
            if (!POOLED) {
                const uint32_t c = covIdx[g];
#pragma unroll
                for (int r = 0; r < FR; ++r) {
                    const int t = min(t0 + 32 * r, T - 1);
                    xx[r]       = 0;
#pragma unroll
                    for (int w4 = 0; w4 < DW / 4; ++w4) {
                        uint4 v = reinterpret_cast<const uint4*>(qx + ((size_t)t * C + c) * DW * 4)[w4];
                        fx[r][4 * w4 + 0] = v.x;
                        fx[r][4 * w4 + 1] = v.y;
                        fx[r][4 * w4 + 2] = v.z;
                        fx[r][4 * w4 + 3] = v.w;
                    }
#pragma unroll
                    for (int w = 0; w < DW; ++w)
                        xx[r] = __dp4a(fx[r][w], fx[r][w], xx[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < FR; ++r) {
                uint32_t dot = 0;
#pragma unroll
                for (int w = 0; w < DW; ++w)
                    dot = __dp4a(mw[w], fx[r][w], dot);
                const int32_t s = (int32_t)(c2 + xx[r] - 2u * dot);
                if (s < minScore[r]) {
                    minScore[r] = s;
                    bestDns[r]  = g - a;
                }
            }
        }
#pragma unroll
        for (int r = 0; r < FR; ++r) {
            const int t = t0 + 32 * r;
            if (t < T) {
                const size_t o = (size_t)t * ld + m;
                if (q)
                    q[o] = minScore[r];
                if (best)
                    best[o] = bestDns[r];
                if (score)
                    score[o] = scalingSquared < 0.0  // batch-int rule (gmm_quant_final_score_rule): one f32 division
                                       ? __fdiv_rn((float)minScore[r], (float)(-2.0 * scalingSquared))
                                       : (float)(0.5 * (double)minScore[r] / scalingSquared);
            }
        }
    }
}

template<int DW, int FR>
int launchScore(gmm_handle* h, uint32_t T, const uint8_t* d_qx, int32_t* d_q, uint32_t* d_best, float* d_score,
                size_t ld) {
    const QuantModel& qm = h->quant;
    const int framesPerBlock = 8 * 32 * FR;
    const int gy             = (T + framesPerBlock - 1) / framesPerBlock;
    int       wantX          = (4 * h->smCount + gy - 1) / gy;
    int       mixPerBlock    = (qm.M + wantX - 1) / (wantX < 1 ? 1 : wantX);
    if (mixPerBlock < 1)
        mixPerBlock = 1;
    const int gx = (qm.M + mixPerBlock - 1) / mixPerBlock;
    dim3      grid(gx, gy), block(256);
    h->timeBegin();
    if (qm.C == 1)
        quant_score_kernel<DW, FR, true><<<grid, block, 0, h->stream>>>(
                d_qx, (int)T, (int)qm.M, (int)qm.C, qm.qmean, qm.covIdx, qm.qconst2, qm.offsets, mixPerBlock,
                qm.finalScoreArg(), d_q, d_best, d_score, ld);
    else
        quant_score_kernel<DW, FR, false><<<grid, block, 0, h->stream>>>(
                d_qx, (int)T, (int)qm.M, (int)qm.C, qm.qmean, qm.covIdx, qm.qconst2, qm.offsets, mixPerBlock,
                qm.finalScoreArg(), d_q, d_best, d_score, ld);
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    return GMM_OK;
}

}  // namespace

int gmm_quant_launch(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best, float* d_score,
                     uint8_t* d_qx, size_t ld) {
    const QuantModel& qm = h->quant;
    if (T == 0)
        return GMM_OK;
    if (h->qtensor.valid)  // integer tensor path (gmm_qtc.cu); this file's dp4a kernel serves every other model
        return gmm_qtc_score(h, T, d_x, d_q, d_best, d_score, d_qx, ld);
    // frames are processed in chunks so that the per-covariance quantized copies stay bounded
    const size_t perFrame = (size_t)qm.C * qm.Dpad;
    uint32_t     chunk    = T;
    if (!d_qx) {
        const size_t cap = (size_t)256 << 20;
        size_t       c   = cap / perFrame;
        if (c < 1)
            c = 1;
        if (c < chunk)
            chunk = (uint32_t)c;
    }
    for (uint32_t t0 = 0; t0 < T; t0 += chunk) {
        const uint32_t n  = (T - t0 < chunk) ? T - t0 : chunk;
        uint8_t*       qx = nullptr;
        if (d_qx)
            qx = d_qx + (size_t)t0 * perFrame;
        else {
            void* p  = nullptr;
            int   rc = h->reserve(SCR_QX, (size_t)n * perFrame, &p);
            if (rc)
                return rc;
            qx = (uint8_t*)p;
        }
        const size_t total  = (size_t)n * qm.C * (qm.Dpad / 4);
        int          blocks = (int)((total + 255) / 256);
        if (blocks > h->smCount * 16)
            blocks = h->smCount * 16;
        quantize_kernel<<<blocks, 256, 0, h->stream>>>(d_x + (size_t)t0 * qm.D, (int)n, (int)qm.D, (int)qm.Dpad,
                                                       (int)qm.C, qm.isrS, qx);
        h->launches++;
        GMM_CUDA_TRY(h, cudaGetLastError());
        if (d_q || d_best || d_score) {
            int32_t*  q  = d_q ? d_q + (size_t)t0 * ld : nullptr;
            uint32_t* b  = d_best ? d_best + (size_t)t0 * ld : nullptr;
            float*    s  = d_score ? d_score + (size_t)t0 * ld : nullptr;
            int       rc = GMM_ERR_UNSUPPORTED;
            switch (qm.Dpad / 4) {
            case 4: rc = launchScore<4, 4>(h, n, qx, q, b, s, ld); break;
            case 8: rc = launchScore<8, 4>(h, n, qx, q, b, s, ld); break;
            case 12: rc = launchScore<12, 4>(h, n, qx, q, b, s, ld); break;
            case 16: rc = launchScore<16, 2>(h, n, qx, q, b, s, ld); break;
            case 20: rc = launchScore<20, 2>(h, n, qx, q, b, s, ld); break;
            case 24: rc = launchScore<24, 2>(h, n, qx, q, b, s, ld); break;
            case 28: rc = launchScore<28, 2>(h, n, qx, q, b, s, ld); break;
            case 32: rc = launchScore<32, 2>(h, n, qx, q, b, s, ld); break;
            default: h->setError("quantized kernel: padded dimension above 128 is not supported"); break;
            }
            if (rc)
                return rc;
        }
    }
    return GMM_OK;
}

// ---- entry point that allocates nothing (no counterpart in the reference,
// see include/gmm_b200.h) ---------------------------------------------------------------------------
// The dp4a scorer above needs one scratch array, the quantized frames qx [T x C x Dpad].  gmm_quant_launch takes it from
// the handle's grow-on-demand scratch (cudaMalloc); this variant keeps it in a file-scope device array instead, so the
// call consists of two kernel launches and nothing else: usable under stream capture and from loops that must not
// call cudaMalloc.  Every GPU that loads the library holds the array (GMM_QUANT_STATIC_SCRATCH_BYTES) until exit.
// All calls on a device share it and it keeps the previous call's bytes: quantize_kernel rewrites every byte the
// scoring kernel reads (padding included), so nothing has to be cleared, but two calls whose kernels overlap on
// the GPU (different streams / handles of one device) race on it.
static __device__ __align__(256) uint8_t g_quantStaticScratch[GMM_QUANT_STATIC_SCRATCH_BYTES];

namespace {

// In host code `g_quantStaticScratch` is the host-side stand-in of the symbol, not a device address: the address a
// kernel may use comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kMaxDevices = 64;
uint8_t*      g_staticScratchAddr[kMaxDevices] = {};
std::mutex    g_staticScratchLock;

int staticScratchSetUp(gmm_handle* h, uint8_t** out) {
    if (h->device < 0 || h->device >= kMaxDevices) {
        h->setError("gmm_quant_score_device_static: device ordinal out of range");
        return GMM_ERR_ARG;
    }
    std::lock_guard<std::mutex> guard(g_staticScratchLock);
    if (!g_staticScratchAddr[h->device]) {
        void* p = nullptr;
        GMM_CUDA_TRY(h, cudaSetDevice(h->device));
        GMM_CUDA_TRY(h, cudaGetSymbolAddress(&p, g_quantStaticScratch));
        if (!p) {
            h->setError("gmm_quant_score_device_static: cudaGetSymbolAddress returned a null address");
            return GMM_ERR_CUDA;
        }
        g_staticScratchAddr[h->device] = (uint8_t*)p;
    }
    *out = g_staticScratchAddr[h->device];
    return GMM_OK;
}

// launchScore without the timing brackets (those create CUDA events when gmm_kernel_timing is on)
template<int DW, int FR>
int launchScoreStatic(gmm_handle* h, uint32_t T, const uint8_t* d_qx, int32_t* d_q, uint32_t* d_best, float* d_score,
                      size_t ld) {
    const QuantModel& qm             = h->quant;
    const int         framesPerBlock = 8 * 32 * FR;
    const int         gy             = (T + framesPerBlock - 1) / framesPerBlock;
    const int         wantX          = std::max(1, (4 * h->smCount + gy - 1) / gy);
    const int         mixPerBlock    = std::max(1, (int)((qm.M + wantX - 1) / wantX));
    dim3              grid((qm.M + mixPerBlock - 1) / mixPerBlock, gy), block(256);
    if (qm.C == 1)
        quant_score_kernel<DW, FR, true><<<grid, block, 0, h->stream>>>(
                d_qx, (int)T, (int)qm.M, (int)qm.C, qm.qmean, qm.covIdx, qm.qconst2, qm.offsets, mixPerBlock,
                qm.finalScoreArg(), d_q, d_best, d_score, ld);
    else
        quant_score_kernel<DW, FR, false><<<grid, block, 0, h->stream>>>(
                d_qx, (int)T, (int)qm.M, (int)qm.C, qm.qmean, qm.covIdx, qm.qconst2, qm.offsets, mixPerBlock,
                qm.finalScoreArg(), d_q, d_best, d_score, ld);
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    return GMM_OK;
}

}  // namespace

extern "C" int gmm_quant_score_device_static(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q,
                                             uint32_t* d_best, float* d_score, size_t ld) {
    if (!h)
        return GMM_ERR_ARG;
    const QuantModel& qm = h->quant;
    if (!qm.valid) {
        h->setError("gmm_quant_score_device_static: no quantized model uploaded");
        return GMM_ERR_ARG;
    }
    if (ld < qm.M || (T && !d_x)) {
        h->setError("gmm_quant_score_device_static: invalid argument (ld < M or NULL features)");
        return GMM_ERR_ARG;
    }
    const size_t need = (size_t)T * qm.C * qm.Dpad;
    if (need > (size_t)GMM_QUANT_STATIC_SCRATCH_BYTES) {
        char msg[256];
        snprintf(msg, sizeof msg,
                 "gmm_quant_score_device_static: %u frames x %u covariances x %u padded dims = %zu bytes of quantized "
                 "frames exceed the static scratch of %zu bytes (at most %zu frames per call for this model)",
                 T, qm.C, qm.Dpad, need, (size_t)GMM_QUANT_STATIC_SCRATCH_BYTES,
                 (size_t)GMM_QUANT_STATIC_SCRATCH_BYTES / ((size_t)qm.C * qm.Dpad));
        h->setError(msg);
        return GMM_ERR_ARG;
    }
    uint8_t* qx = nullptr;
    int      rc = staticScratchSetUp(h, &qx);
    if (rc)
        return rc;
    if (T == 0)
        return GMM_OK;
    GMM_CUDA_TRY(h, cudaSetDevice(h->device));
    const size_t total  = (size_t)T * qm.C * (qm.Dpad / 4);
    const int    blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)h->smCount * 16);
    quantize_kernel<<<blocks, 256, 0, h->stream>>>(d_x, (int)T, (int)qm.D, (int)qm.Dpad, (int)qm.C, qm.isrS, qx);
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    if (!d_q && !d_best && !d_score)
        return GMM_OK;
    switch (qm.Dpad / 4) {
    case 4: return launchScoreStatic<4, 4>(h, T, qx, d_q, d_best, d_score, ld);
    case 8: return launchScoreStatic<8, 4>(h, T, qx, d_q, d_best, d_score, ld);
    case 12: return launchScoreStatic<12, 4>(h, T, qx, d_q, d_best, d_score, ld);
    case 16: return launchScoreStatic<16, 2>(h, T, qx, d_q, d_best, d_score, ld);
    case 20: return launchScoreStatic<20, 2>(h, T, qx, d_q, d_best, d_score, ld);
    case 24: return launchScoreStatic<24, 2>(h, T, qx, d_q, d_best, d_score, ld);
    case 28: return launchScoreStatic<28, 2>(h, T, qx, d_q, d_best, d_score, ld);
    case 32: return launchScoreStatic<32, 2>(h, T, qx, d_q, d_best, d_score, ld);
    default: break;
    }
    h->setError("gmm_quant_score_device_static: padded dimension above 128 is not supported");
    return GMM_ERR_UNSUPPORTED;
}
