// Quantized (u8) scorer on the integer tensor path of sm_100a (tcgen05.mma kind::i8, u8 x u8 -> s32, exact):
// bit-identical to the reference's SIMD-diagonal-maximum scorer
// (/root/reference/src/Mm/SimdFeatureScorer.cc:135-176, IntelOptimization.cc:50-113) for pooled-covariance models;
// everything else stays on the dp4a kernel of gmm_quant.cu, which is the same arithmetic on CUDA cores.
//
//   q[t][m] = min_k( qconst_k + sum_d (qmean_k[d] - qx_t[d])^2 )              (first strict minimum)
//           = min_k( c2_k - 2 dot(qmean_k, qx_t) ) + |qx_t|^2,                  c2_k = qconst_k + |qmean_k|^2
// The contraction dot() is one dense u8 GEMM: frames x Gaussians, accumulated in s32 in TMEM.  Every partial is an
// exact integer, so nothing is approximated.
//
// No per-column arithmetic in the epilogue: the tensor core adds the per-Gaussian constant itself.  The operand rows
// are K = Dpad + S bytes long (S = 16 or 32 spare bytes); the frames carry 255, ..., 255, 1 there and column k carries
// the base-255 digits of a constant, so the contraction over the spare bytes yields that constant exactly.  The k-steps
// are issued TWICE into the same accumulator, the second time with a variant of the last k-step whose spare bytes hold
// a second constant:
//   acc_k = 2 dot_k + E_k,   E_k = E1_k + E2_k = cmax - c2_k >= 0   (cmax = max_k c2_k over the model)
//   s_k   = c2_k - 2 dot_k = cmax - acc_k,        q = (cmax + |qx|^2) - max_k acc_k
// so the maximum approximation is a bare unsigned max tree over the TMEM values (VIMNMX3), and the reference's first
// strict minimum (lowest density index among equal scores) is the max over  key_k = (acc_k << 8) | (255 - slot_k):
// one integer multiply-add per element, needed only when the arg-min is asked for.  On tiles whose mixtures all have the
// same shape the (255 - slot) term is a compile-time immediate; other tiles read it from a table shipped with the tile.
// Requirements checked at packing time (else the model is left to the dp4a kernel): one covariance, mixtures of at most
// 128 densities, K <= 128, cmax - cmin within what two sets of spare bytes can hold (2 * (255 * 255 * (S - 1) + 254)),
// 2 Dpad 255^2 + (cmax - cmin) < 2^24 (the key fits 32 bits; nothing wraps: zero-weight densities, whose reference
// constant is the x86 "integer indefinite" 0x80000000, fail this and take the exact wrapping path of gmm_quant.cu).
//
// Layout and pipeline are those of the float tensor kernel (gmm_tc.cu): pre-swizzled tile images in HBM staged
// by cp.async.bulk, B tile resident per work item, A tiles streaming through a ring, two accumulators in TMEM,
// tile-indexed ready/done barrier rings, one MMA-issuing lane, eight or sixteen epilogue warps.
//   B tile = 256 packed columns: [atoms][256 rows][64 B] u8 (k-steps 0 .. kS-1, then the variant of k-step kS-1)
//                                | Kc[256] u32 (255 - slot; 0 for a pad column) | TileMeta
//   A tile = 128 frames:         [atoms][128 rows][64 B] u8          + per-frame |qx|^2 (u32)
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <type_traits>

#include "gmm_internal.cuh"
#include "gmm_umma.cuh"

namespace {

constexpr int      kQSwz       = 64;                   // swizzle atom: 64-byte rows hold two 32-byte k-steps
constexpr int      kQMetaBytes = kTileN * 4 + 48;      // Kc[256] u32, TileMeta
constexpr uint32_t kNoKey      = 0u;                   // pad / filler column (zero operand row, Kc = 0): never the maximum

struct QKParams {
    const uint8_t*  bTiles;
    const uint8_t*  aTiles;
    const uint32_t* rowXX;  // per frame row: |qx|^2
    int32_t*        q;
    uint32_t*       best;
    float*          score;
    size_t          ld;
    double          scalingSquared;
    float           twoScalingSquared;  // f32 path of the final score (OUT & 4): q / (2 s^2), see qtc_launch
    uint32_t        T;
    uint32_t        nTiles, frameTiles, split;
    uint32_t        swapTile;    // B tile that trades places with the last one (nTiles - 1: none), see qtc_launch
    uint32_t        kSteps;      // 32-byte k-steps of the contraction (the B image holds one more: the variant)
    uint32_t        atoms;       // 64-byte swizzle atoms per A operand row
    uint32_t        atomsB;      // ... per B operand row (k-steps 0 .. kSteps, the last being the variant)
    uint32_t        aTileBytes, bTileBytes;
    uint32_t        stagesA, buffersB;
    uint32_t        stagePitch;  // words per staging row (maxMixPerPart + 1)
    uint32_t        qBase;       // cmax (wrapping u32): q = qBase + |qx|^2 - max_k acc_k
    unsigned long long* dbg;     // GMM_QTC_DEBUG: per-CTA cycle counters (16 words per CTA)
    uint32_t        vecStores;   // 0: always the scalar staged store path (GMM_QTC_STORES=scalar)
    uint32_t        dbgSkip;     // GMM_QTC_DEBUG_SKIP (timing experiments only, results are NOT written): 1 = the epilogue
                                 // only hands the accumulator back, 2 = it loads and reduces but stores nothing
    uint32_t        keyMul;      // 256, as a run-time value: key = acc * keyMul + term compiles to IMAD (FMA pipe)
};

__device__ __forceinline__ uint8_t quantize_u8(float v) {
    // round(): half away from zero; (int) of an out-of-range value is cvttss2si's 0x80000000
    // (/root/reference/src/Mm/Utilities.hh:179-191)
    float r = roundf(v);
    int   i = (fabsf(r) < 2147483648.f) ? __float2int_rz(r) : (int)0x80000000;
    int   s = (int)((unsigned)i + 128u);
    s       = s < 0 ? 0 : (s > 255 ? 255 : s);
    return (uint8_t)s;
}

// Frame preparation: x [T x D] -> A tile images (quantized bytes, swizzled; spare bytes [Dpad, K) = 255, ..., 255, 1,
// the multipliers of the constants' digits in the B rows) + |qx|^2 per frame; optionally the plain quantized frames
// [T x Dpad] for the caller.  Eight lanes share a frame, lane c owns bytes [16c, 16c + 16) (K <= 128).
__global__ void __launch_bounds__(256) qtc_prep_kernel(const float* __restrict__ x, uint32_t T, uint32_t D,
                                                       uint32_t Dpad, uint32_t K, uint32_t frameTiles,
                                                       const float* __restrict__ isrS, uint32_t atoms,
                                                       uint8_t* __restrict__ aTiles, uint32_t* __restrict__ rowXX,
                                                       uint8_t* __restrict__ qxOut) {
    const uint32_t lane = threadIdx.x & 31, c = lane & 7;
    const uint32_t rows = frameTiles * kTileM, chunksAll = atoms * (kQSwz / 16);
    const uint32_t framesPerGrid = gridDim.x * (blockDim.x >> 5) * 4;
    for (uint32_t t = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 4 + (lane >> 3); t < rows;
         t += framesPerGrid) {
        uint32_t w[4] = {0, 0, 0, 0};
        if (t < T && c * 16 < Dpad) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const uint32_t d = c * 16 + j;
                uint32_t       v = 0;
                if (d < D)
                    v = quantize_u8(__fmul_rn(x[(size_t)t * D + d], isrS[d]));
                w[j >> 2] |= v << (8 * (j & 3));
            }
        }
        uint32_t xx = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            xx = __dp4a(w[i], w[i], xx);
#pragma unroll
        for (int o = 4; o > 0; o >>= 1)
            xx += __shfl_xor_sync(0xffffffffu, xx, o);
        if (c == 0)
            rowXX[t] = xx;
        if (qxOut && t < T && c * 16 < Dpad)
            *reinterpret_cast<uint4*>(qxOut + (size_t)t * Dpad + c * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        if (c * 16 >= Dpad && c * 16 < K) {  // spare bytes (rows past T as well: their results are never stored)
            w[0] = w[1] = w[2] = 0xFFFFFFFFu;
            w[3] = (c * 16 + 16 == K) ? 0x01FFFFFFu : 0xFFFFFFFFu;
        }
        const uint32_t tile = t / kTileM, row = t % kTileM;
        uint8_t*       base = aTiles + (size_t)tile * atoms * kTileM * kQSwz;
        if (c < chunksAll)  // bytes past K inside the last atom are written as zeros (w == 0)
            *reinterpret_cast<uint4*>(base + (size_t)(c / 4) * kTileM * kQSwz + swz_offset(row, c % 4, kQSwz)) =
                    make_uint4(w[0], w[1], w[2], w[3]);
    }
}

__device__ __forceinline__ void tc_mma_i8(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(dTmem),
            "l"(aDesc), "l"(bDesc), "r"(idesc), "r"(accumulate)
            : "memory");
}

struct QEpiState {
    uint32_t run;  // running maximum (key or raw accumulator) of the open mixture
    uint32_t sp;   // shared-memory address of this row's next staged q (the arg-min plane lies offB behind)
};

// One 32-column chunk (four 8-column segments) of a row's s32 accumulator (acc = 2 dot + E >= 0).
// SP > 0: every mixture of the column part spans exactly SP segments and none is empty (TileMeta::uni): closing
// positions are compile-time constants, and
//   !WB: the reduction is a max tree over the raw TMEM values, nothing else;
//    WB: key = (acc << 8) + (255 - slot) with the slot term an immediate.
// SP = 0: key = (acc << 8) + Kc with Kc read from the tile's table; segEnd bits close the mixtures; a mixture without
// densities (all its columns pads: key 0) yields the reference's INT_MAX / no density.
// resQ != nullptr (SP > 0 only): the closed mixtures' q (and arg-min, resB) stay in registers, index seg / SP, for the
// vector store path instead of being staged one by one.
// The epilogue is bound by instruction issue on the two half-rate integer pipes: shifts, LEA and the three-input maxima
// go to the ALU pipe, IMAD to the FMA pipe.  Three of four keys are therefore built by a multiply-add whose factor 256
// arrives as a run-time value (ptxas turns a literal 256 back into a shift): 0.5 + 0.25 ALU and 0.75 FMA instructions
// per column instead of 1.5 ALU.
__device__ __forceinline__ uint32_t qkey(uint32_t v, uint32_t term, uint32_t mul, bool viaMad) {
    return viaMad ? v * mul + term : (v << 8) + term;
}

template<int SP, bool WB, int MAD = 3>
__device__ __forceinline__ void qepi_chunk(const uint32_t* v, int ch, uint32_t kcAddr, uint32_t segEnd, uint32_t qbx,
                                           uint32_t offB, QEpiState& st, uint32_t mul, uint32_t* resQ = nullptr,
                                           uint32_t* resB = nullptr) {
    uint32_t kc[32];
    if (SP == 0) {  // the chunk's 32 per-column terms first (eight 128-bit loads in flight together)
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(kc[4 * i]), "=r"(kc[4 * i + 1]), "=r"(kc[4 * i + 2]), "=r"(kc[4 * i + 3])
                         : "r"(kcAddr + ch * 128 + i * 16));
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int seg = ch * 4 + s;
        uint32_t  key[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (SP > 0 && !WB)
                key[j] = v[s * 8 + j];
            else if (SP > 0)
                key[j] = qkey(v[s * 8 + j], (uint32_t)(255 - ((seg % SP) * 8 + j)), mul, (j & 3) < MAD);
            else
                key[j] = qkey(v[s * 8 + j], kc[s * 8 + j], mul, (j & 3) < MAD);
        }
        // eight keys and the running maximum of the open mixture: four three-input maxima
        const bool     first = SP > 0 && (seg % SP) == 0;
        const uint32_t m     = __vimax3_u32(__vimax3_u32(key[0], key[1], key[2]), __vimax3_u32(key[3], key[4], key[5]),
                                            first ? max(key[6], key[7]) : __vimax3_u32(key[6], key[7], st.run));
        bool close;
        if (SP > 0)
            close = (seg % SP) == SP - 1;
        else
            close = (segEnd >> seg) & 1u;
        if (close) {  // uniform across the warp
            if (SP > 0 && resQ) {
                resQ[seg / SP] = WB ? qbx - (m >> 8) : qbx - m;
                if (WB)
                    resB[seg / SP] = 255u - (m & 255u);
            }
            else if (SP > 0 && !WB)
                sts_u32(st.sp, qbx - m);
            else {
                const bool none = SP == 0 && m == kNoKey;
                sts_u32(st.sp, none ? 0x7FFFFFFFu : qbx - (m >> 8));
                if (WB)
                    sts_u32(st.sp + offB, none ? 0xFFFFFFFFu : 255u - (m & 255u));
            }
            st.sp += 4;
            st.run = kNoKey;
        }
        else
            st.run = m;
    }
}

// Persistent kernel: warp 0 bulk-copy producer (+ TMEM allocation), warp 1 MMA issuer, NE = 8 or 16 epilogue warps
// (TMEM lane quarter = warp & 3, column part of 1024 / NE columns = (warp - 2) / 4; 16 warps when no mixture has more
// than 64 densities: the epilogue, not the two MMAs per tile, bounds this kernel).  Barrier protocol as in gmm_tc.cu.
// OUT: which arrays the caller asked for, fixed at compile time so that the store loop is branch-free -- bit 0 q, bit 1
// arg-min, bit 2 final f32 score by ONE f32 division (exact, see qtc_launch); OUT = 0 keeps all three as run-time tests
// and the score as the reference's f64 expression (any other combination, or |q| beyond 2^24).
// NG = 2 (NE = 8 only): TWO epilogue groups of eight warps; group g drains the accumulator of the tiles with t % 2 == g,
// so one group's reduction (integer pipes) runs while the other group's row pieces leave (load/store unit, memory
// latency) -- with one group the two phases alternate and each pipe idles half the time.
template<int NE, int NG, int OUT, int MAD = 3>
__global__ void __launch_bounds__(32 * (kEpiWarp0 + NE * NG), 1) qtc_score_kernel(const QKParams p) {
    static_assert(NG == 1 || (NG == 2 && NE == 8 && kRing % 2 == 0), "two groups: eight warps each, ring index parity = tile parity");
    constexpr bool WQ = OUT == 0 || (OUT & 1), WB = OUT == 0 || (OUT & 2), WS = OUT == 0 || (OUT & 4);
    extern __shared__ uint8_t smemRaw[];
    uint8_t*  smem = smemRaw + ((1024u - (smem_u32(smemRaw) & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int PW = 1024 / NE, SEGS = PW / 8, SWZ = kQSwz;

    // carve shared memory: [B buffers][A stages][staging: q plane, arg-min plane][barriers]
    const uint32_t bBytes    = (p.bTileBytes + 1023u) & ~1023u;
    uint8_t*       sB        = smem;
    uint8_t*       sA        = sB + (size_t)p.buffersB * bBytes;
    uint32_t*      sStage    = reinterpret_cast<uint32_t*>(sA + (size_t)p.stagesA * p.aTileBytes);
    // staging: one [32 rows x stagePitch] block per store group and plane.  8 warps: a group is one warp.  16 warps:
    // the four warps of a TMEM lane quarter stage the 64-column parts of their 32 rows side by side (two blocks,
    // alternating by tile, so one named barrier per tile is enough) and store the rows together: pieces of a whole
    // tile's mixtures (64 bytes per row at 16 densities per mixture), not of a part's (16 bytes)
    constexpr int  kGroups = NE == 16 ? 4 : 8, kBufs = NE == 16 ? 2 : 1, GW = NE == 16 ? 128 : 32;
    const uint32_t stageWarp = 32 * p.stagePitch;  // words per block
    uint64_t*      bars      = reinterpret_cast<uint64_t*>(
            (reinterpret_cast<uintptr_t>(sStage + 2 * kGroups * kBufs * NG * stageWarp) + 15) & ~(uintptr_t)15);
    uint64_t* ready    = bars;              // [kRing] tile t may be issued (A landed, accumulator drained)
    uint64_t* done     = bars + kRing;      // [kRing] the MMAs of tile t have retired
    uint64_t* fullB    = bars + 2 * kRing;  // [2]
    uint64_t* emptyB   = fullB + 2;         // [2]
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(emptyB + 2);

    if (threadIdx.x == 32) {
        for (int i = 0; i < kRing; ++i) {
            mbar_init(smem_u32(&ready[i]), 1 + NE);
            mbar_init(smem_u32(&done[i]), 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(smem_u32(&fullB[i]), 1);
            mbar_init(smem_u32(&emptyB[i]), 1 + NE * NG);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmemSlot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmemBase = *tmemSlot;

    const uint32_t items        = p.nTiles * p.split;
    const uint32_t tilesPerItem = (p.frameTiles + p.split - 1) / p.split;

    if (warp == 0) {
        // ===== producer =====
        if (lane == 0) {
            uint32_t itemCount = 0, t = 0, stage = 0;
            for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
                const uint32_t nt0 = item / p.split, chunk = item % p.split;
                const uint32_t nt  = nt0 == p.nTiles - 1 ? p.swapTile : nt0 == p.swapTile ? p.nTiles - 1 : nt0;
                const uint32_t f0 = chunk * tilesPerItem, f1 = min(p.frameTiles, f0 + tilesPerItem);
                if (f0 >= f1)
                    continue;
                const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
                mbar_wait(smem_u32(&emptyB[b]), (useB & 1) ^ 1);
                mbar_expect_tx(smem_u32(&fullB[b]), p.bTileBytes);
                bulk_g2s(smem_u32(sB + (size_t)b * bBytes), p.bTiles + (size_t)nt * p.bTileBytes, p.bTileBytes,
                         smem_u32(&fullB[b]));
                for (uint32_t f = f0; f < f1; ++f, ++t) {
                    if (t >= p.stagesA) {
                        const uint32_t u = t - p.stagesA;
                        mbar_wait(smem_u32(&done[u % kRing]), (u / kRing) & 1);
                    }
                    const uint32_t rb = smem_u32(&ready[t % kRing]);
                    mbar_expect_tx(rb, p.aTileBytes);
                    bulk_g2s(smem_u32(sA + (size_t)stage * p.aTileBytes), p.aTiles + (size_t)f * p.aTileBytes,
                             p.aTileBytes, rb);
                    if (++stage == p.stagesA)
                        stage = 0;
                }
            }
        }
    }
    else if (warp == 1) {
        // ===== MMA issuer (warp converged, one elected lane issues) =====
        // instruction descriptor: D = s32, A = B = u8, K-major both, N = 256, M = 128
        constexpr uint32_t idesc    = (2u << 4) | ((uint32_t)(kTileN >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
        constexpr uint32_t kPerAtom = SWZ / 32;
        constexpr uint32_t descHi   = (uint32_t)((8 * SWZ) >> 4) | (1u << 14) | (4u << 29);  // SBO, version, SWIZZLE_64B
        const bool         leader   = elect_one();
        const uint32_t     kS       = p.kSteps;
        uint32_t           itemCount = 0, t = 0, stage = 0;
        bool               have = false;
        long long          mW = 0, mI = 0, mc0 = 0;
        const long long    mStart = clock64();
        unsigned long long gt0 = 0;
        if (p.dbg) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt0));
        for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
            const uint32_t chunk = item % p.split;
            const uint32_t f0 = chunk * tilesPerItem, f1 = min(p.frameTiles, f0 + tilesPerItem);
            if (f0 >= f1)
                continue;
            const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
            mbar_wait(smem_u32(&fullB[b]), useB & 1);
            const uint32_t bLo = (smem_u32(sB + (size_t)b * bBytes) & 0x3FFFF) >> 4;
            for (uint32_t f = f0; f < f1; ++f, ++t) {
                if (p.dbg) mc0 = clock64();
                if (!have)
                    mbar_wait(smem_u32(&ready[t % kRing]), (t / kRing) & 1);
                tc_fence_after();
                if (p.dbg) { const long long c1 = clock64(); mW += c1 - mc0; mc0 = c1; }
                const uint32_t aLo   = (smem_u32(sA + (size_t)stage * p.aTileBytes) & 0x3FFFF) >> 4;
                const uint32_t dTmem = tmemBase + (t & 1) * kTileN;
                if (leader) {
                    // every k-step twice (acc = 2 dot + E1 + E2): the second pass reads the variant of the last k-step,
                    // stored behind the others in the B image, whose spare bytes hold the second constant
                    for (uint32_t i = 0; i < 2 * kS; ++i) {
                        const uint32_t ks = i < kS ? i : i - kS, kb = (i == 2 * kS - 1) ? kS : ks;
                        const uint32_t ao = aLo + (ks / kPerAtom) * (kTileM * SWZ / 16) + (ks % kPerAtom) * 2;
                        const uint32_t bo = bLo + (kb / kPerAtom) * (kTileN * SWZ / 16) + (kb % kPerAtom) * 2;
                        tc_mma_i8(dTmem, ((uint64_t)descHi << 32) | ao, ((uint64_t)descHi << 32) | bo, idesc, i ? 1u : 0u);
                    }
                    tc_commit(smem_u32(&done[t % kRing]));
                    if (f + 1 == f1)
                        tc_commit(smem_u32(&emptyB[b]));
                }
                have = mbar_test(smem_u32(&ready[(t + 1) % kRing]), ((t + 1) / kRing) & 1);
                __syncwarp();
                if (++stage == p.stagesA)
                    stage = 0;
                if (p.dbg) mI += clock64() - mc0;
            }
        }
        if (p.dbg && leader) {
            unsigned long long* d = p.dbg + blockIdx.x * 128;
            unsigned long long gt1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt1));
            uint32_t smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            d[0] = clock64() - mStart, d[1] = mW, d[2] = mI, d[4] = t, d[5] = gt0, d[6] = gt1, d[7] = smid;
        }
    }
    else {
        // ===== epilogue: TMEM -> registers -> packed-key min per mixture -> staged -> global =====
        const int      grp = (warp - kEpiWarp0) / NE, e = (warp - kEpiWarp0) % NE, qd = warp & 3, part = e >> 2;
        const uint32_t gi     = NE == 16 ? (uint32_t)qd : (uint32_t)e;       // store group
        const uint32_t gl     = NE == 16 ? (uint32_t)(part * 32 + lane) : (uint32_t)lane;  // thread within the group
        const uint32_t stGrp  = smem_u32(sStage) + ((uint32_t)grp * 2 * kGroups + gi) * kBufs * stageWarp * 4;
        const uint32_t offB   = kGroups * kBufs * stageWarp * 4;
        // the three outputs share one indexing: address everything relative to a base array that is present
        char* const     gBase = reinterpret_cast<char*>(p.q ? (void*)p.q : p.best ? (void*)p.best : (void*)p.score);
        const ptrdiff_t bestDelta  = p.best ? reinterpret_cast<char*>(p.best) - gBase : 0;
        const ptrdiff_t scoreDelta = p.score ? reinterpret_cast<char*>(p.score) - gBase : 0;
        uint32_t       itemCount = 0, t = 0;
        long long      eW = 0, eL = 0, eB = 0, eS = 0, ec0 = 0, eI = 0;
        const long long eStart = clock64();
        if (lane == 0) {  // both accumulators start drained (two groups: each speaks for its own)
            if (NG == 1 || grp == 0)
                mbar_arrive(smem_u32(&ready[0]));
            if (NG == 1 || grp == 1)
                mbar_arrive(smem_u32(&ready[1]));
        }
        for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
            const uint32_t chunk = item % p.split;
            const uint32_t f0 = chunk * tilesPerItem, f1 = min(p.frameTiles, f0 + tilesPerItem);
            if (f0 >= f1)
                continue;
            const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
            if (p.dbg) ec0 = clock64();
            mbar_wait(smem_u32(&fullB[b]), useB & 1);
            if (p.dbg) eI += clock64() - ec0;
            // this group's first tile of the item (tiles alternate between the groups across items as well)
            const uint32_t  fG = NG == 2 ? f0 + ((t ^ (uint32_t)grp) & 1u) : f0;
            const uint8_t*  tile = sB + (size_t)b * bBytes + (size_t)p.atomsB * kTileN * SWZ;
            const uint32_t  kcAddr = smem_u32(tile) + (uint32_t)part * PW * 4;  // Kc of this warp's columns
            const uint32_t* meta = reinterpret_cast<const uint32_t*>(tile + kTileN * 4);  // TileMeta words
            constexpr uint32_t segMask = (1u << SEGS) - 1u;
            const uint32_t  segEnd  = (meta[0] >> (SEGS * part)) & segMask;
            const uint32_t  nMixW   = __popc(segEnd);  // mixtures this warp closes
            // the store group's mixtures: this warp's (8 warps), or the whole tile's (16 warps; colOff = where this
            // warp's mixtures start among them)
            uint32_t nMix = nMixW, colOff = 0, mixBase = meta[2 + part];
            if (NE == 16) {
                const uint32_t all = meta[0];
                nMix               = __popc(all);
                colOff             = __popc(all & ((1u << (SEGS * part)) - 1u));
                mixBase            = 0;
                for (int pp = 3; pp >= 0; --pp)  // mixtures of a tile are consecutive: any part that has some gives the base
                    if ((all >> (SEGS * pp)) & segMask)
                        mixBase = meta[2 + pp] - __popc(all & ((1u << (SEGS * pp)) - 1u));
            }
            const uint32_t  adv_r = GW / max(nMix, 1u), adv_j = GW - adv_r * nMix;
            const uint32_t  lr0 = gl / max(nMix, 1u), lj0 = gl - lr0 * nMix;
            const size_t    gStep = ((size_t)adv_r * p.ld + adv_j) * 4, gStepW = gStep + (p.ld - nMix) * 4;  // bytes
            const uint32_t  sStep = (adv_r * p.stagePitch + adv_j) * 4, sStepW = sStep + (p.stagePitch - nMix) * 4;
            const uint32_t  ssOff = (lr0 * p.stagePitch + lj0) * 4;
            const uint32_t  stOff = ((uint32_t)lane * p.stagePitch + colOff) * 4;
            const uint32_t  passes = (32 * nMix + GW - 1) / GW;
            size_t          gTile = ((size_t)(fG * kTileM + qd * 32 + lr0) * p.ld + mixBase + lj0) * 4;
            const size_t    gTileStep = (size_t)NG * kTileM * p.ld * 4;
            const uint32_t* xxp = p.rowXX + (size_t)fG * kTileM + qd * 32 + lane;
            // VEC (uniform tile, 16-byte aligned row pieces): a lane keeps its R = SEGS / SP results in registers, stages
            // them with 128-bit stores into rows of W + 4 words (W = the group's results per row), and the group writes
            // the block out with one 128-bit load and store per thread and array and pass -- a handful of instructions
            // where the scalar path below spends a hundred (the epilogue is bound by the half-rate integer pipes)
            const uint32_t tItem = t;
            auto tiles = [&](auto spTag, auto vecTag) {
                constexpr int  SP  = decltype(spTag)::value;
                constexpr bool VEC = decltype(vecTag)::value && SP > 0 && (SEGS / (SP > 0 ? SP : 1)) % 4 == 0;
                constexpr int  R   = SP > 0 ? SEGS / SP : 1;          // results per lane and tile
                constexpr int  W   = NE == 16 ? 4 * R : R;            // results per row of the store group
                constexpr int  PV  = W + 4;                           // staging pitch of the vector path (words)
                constexpr int  WQ4 = W / 4 > 0 ? W / 4 : 1;           // 16-byte pieces per row
                uint32_t       resQ[VEC ? R : 1], resB[VEC ? R : 1];
                uint32_t* const rq = VEC ? resQ : nullptr;
                uint32_t* const rb = VEC ? resB : nullptr;
                // vector path: this thread's piece and first row, its global offset (bytes) and the staging offsets
                const uint32_t vqi = gl % WQ4, vr0 = gl / WQ4;
                size_t         gVec = ((size_t)(fG * kTileM + qd * 32 + vr0) * p.ld + mixBase + vqi * 4) * 4;
                const uint32_t vRead  = (vr0 * PV + vqi * 4) * 4;
                const uint32_t vWrite = ((uint32_t)lane * PV + (NE == 16 ? (uint32_t)part * R : 0u)) * 4;
                uint32_t xxNext = fG < f1 ? *xxp : 0u;  // |qx|^2 of the next tile's row, fetched one tile ahead
                for (uint32_t f = fG; f < f1; f += NG) {
                    const uint32_t t      = tItem + (f - f0);  // tile counter of the CTA (shadows the item-level one)
                    const uint32_t acc    = t & 1;
                    const uint32_t xxBase = xxNext + p.qBase;
                    xxp += NG * kTileM;
                    if (f + NG < f1)
                        xxNext = *xxp;
                    if (p.dbg) ec0 = clock64();
                    mbar_wait(smem_u32(&done[t % kRing]), (t / kRing) & 1);
                    tc_fence_after();
                    if (p.dbg) { const long long c1 = clock64(); eW += c1 - ec0; ec0 = c1; }
                    const uint32_t taddr   = tmemBase + ((uint32_t)(qd * 32) << 16) + acc * kTileN + part * PW;
                    const uint32_t drained = smem_u32(&ready[(t + 2) % kRing]);
                    if (p.dbgSkip == 1) {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0)
                            mbar_arrive(drained);
                        continue;
                    }
                    const uint32_t stBuf = stGrp + (kBufs == 2 ? (t & 1) * stageWarp * 4 : 0u);
                    QEpiState      st;
                    st.run = kNoKey, st.sp = stBuf + stOff;  // SP > 0 never reads run before writing it
                    uint32_t va[32];
                    if (NE == 16 && SP > 0) {
                        // no per-column table: both 32-column chunks in flight, the accumulator goes straight back
                        uint32_t vb[32];
                        tmem_ld32(taddr, va);
                        tmem_ld32(taddr + 32, vb);
                        tmem_ld_wait();
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0)
                            mbar_arrive(drained);
                        qepi_chunk<SP, WB, MAD>(va, 0, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                        qepi_chunk<SP, WB, MAD>(vb, 1, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                    }
                    else if (NE == 16) {
                        // 96 registers per thread: one 32-column chunk (+ its 32 table entries) live at a time
                        tmem_ld32(taddr, va);
                        tmem_ld_wait();
                        qepi_chunk<SP, WB, MAD>(va, 0, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                        tmem_ld32(taddr + 32, va);
                        tmem_ld_wait();
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0)
                            mbar_arrive(drained);
                        qepi_chunk<SP, WB, MAD>(va, 1, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                    }
                    else if (NG == 2) {
                        // 96 registers per thread (18 warps): one 32-column chunk live at a time; the other warps of the
                        // scheduler (its second group among them) cover the TMEM load latency
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch) {
                            tmem_ld32(taddr + 32 * ch, va);
                            tmem_ld_wait();
                            if (ch == 3) {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0)
                                    mbar_arrive(drained);
                            }
                            qepi_chunk<SP, WB, MAD>(va, ch, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                        }
                    }
                    else {
                        uint32_t vb[32];
                        tmem_ld32(taddr, va);
                        tmem_ld_wait();
                        tmem_ld32(taddr + 32, vb);
                        qepi_chunk<SP, WB, MAD>(va, 0, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                        tmem_ld_wait();
                        tmem_ld32(taddr + 64, va);
                        qepi_chunk<SP, WB, MAD>(vb, 1, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                        tmem_ld_wait();
                        tmem_ld32(taddr + 96, vb);
                        qepi_chunk<SP, WB, MAD>(va, 2, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                        tmem_ld_wait();
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0)
                            mbar_arrive(drained);
                        qepi_chunk<SP, WB, MAD>(vb, 3, kcAddr, segEnd, xxBase, offB, st, p.keyMul, rq, rb);
                    }
                    if (VEC) {
#pragma unroll
                        for (int i = 0; i < R / 4; ++i) {
                            sts_v4(stBuf + vWrite + i * 16, resQ[4 * i], resQ[4 * i + 1], resQ[4 * i + 2], resQ[4 * i + 3]);
                            if (WB)
                                sts_v4(stBuf + offB + vWrite + i * 16, resB[4 * i], resB[4 * i + 1], resB[4 * i + 2],
                                       resB[4 * i + 3]);
                        }
                    }
                    if (p.dbg) { const long long c1 = clock64(); eL += c1 - ec0; ec0 = c1; }
                    if (p.dbgSkip == 2)
                        continue;
                    if (NE == 16)  // the quarter's four warps have staged their parts of these 32 rows
                        asm volatile("bar.sync %0, 128;" ::"r"(1 + qd) : "memory");
                    else
                        __syncwarp();
                    if (p.dbg) { const long long c1 = clock64(); eB += c1 - ec0; ec0 = c1; }
                    const uint32_t rbase = f * kTileM + qd * 32;
                    const uint32_t rows  = min(32u, p.T > rbase ? p.T - rbase : 0u);
                    if (VEC) {
                        constexpr int RPP = GW / WQ4;  // rows per pass
                        char*         g   = gBase + gVec;
                        gVec += gTileStep;
#pragma unroll
                        for (int ps = 0; ps < 32 / RPP; ++ps) {
                            if (vr0 + ps * RPP < rows) {
                                const uint4 qv = lds_v4(stBuf + vRead + ps * RPP * PV * 4);
                                char*       gp = g + (size_t)ps * RPP * p.ld * 4;
                                if (WQ)
                                    *reinterpret_cast<uint4*>(gp) = qv;
                                if (WB)
                                    *reinterpret_cast<uint4*>(gp + bestDelta) = lds_v4(stBuf + offB + vRead + ps * RPP * PV * 4);
                                if (WS)
                                    *reinterpret_cast<float4*>(gp + scoreDelta) =
                                            make_float4(__fdiv_rn((float)(int32_t)qv.x, p.twoScalingSquared),
                                                        __fdiv_rn((float)(int32_t)qv.y, p.twoScalingSquared),
                                                        __fdiv_rn((float)(int32_t)qv.z, p.twoScalingSquared),
                                                        __fdiv_rn((float)(int32_t)qv.w, p.twoScalingSquared));
                            }
                        }
                        if (NE != 16)
                            __syncwarp();
                        if (p.dbg) eS += clock64() - ec0;
                        continue;
                    }
                    // coalesced store of the staged [32 rows x nMix] block: passes of GW consecutive elements
                    char*          g     = gBase + gTile;  // byte address of this thread's first element in q[]
                    uint32_t       ss    = stBuf + ssOff;
                    gTile += gTileStep;
                    auto put = [&](uint32_t r) {
                        if (r < rows) {
                            const uint32_t qv = lds_u32(ss);
                            if (OUT == 0 ? p.q != nullptr : WQ)
                                *reinterpret_cast<uint32_t*>(g) = qv;
                            if (OUT == 0 ? p.best != nullptr : WB)
                                *reinterpret_cast<uint32_t*>(g + bestDelta) = lds_u32(ss + offB);
                            if (OUT == 0) {
                                if (p.score)  // f32( 0.5 * q / scalingSquared ) in double, SimdFeatureScorer.cc:135-145
                                    *reinterpret_cast<float*>(g + scoreDelta) =
                                            p.scalingSquared < 0.0 ? __fdiv_rn((float)(int32_t)qv, p.twoScalingSquared)
                                                                   : (float)(0.5 * (double)(int32_t)qv / p.scalingSquared);
                            }
                            else if (WS)  // the same bits from one f32 division (|q| < 2^24, see qtc_launch)
                                *reinterpret_cast<float*>(g + scoreDelta) = __fdiv_rn((float)(int32_t)qv, p.twoScalingSquared);
                        }
                    };
                    if (adv_j == 0) {  // nMix divides 32: a lane keeps its column
#pragma unroll 4
                        for (uint32_t i = 0, r = lr0; i < passes; ++i, r += adv_r) {
                            put(r);
                            g += gStep, ss += sStep;
                        }
                    }
                    else {
                        uint32_t r = lr0, j = lj0;
                        for (uint32_t i = 0; i < passes; ++i) {
                            put(r);
                            const bool w = j + adv_j >= nMix;
                            j            = w ? j + adv_j - nMix : j + adv_j;
                            r += adv_r + (w ? 1u : 0u);
                            g += w ? gStepW : gStep, ss += w ? sStepW : sStep;
                        }
                    }
                    if (NE != 16)
                        __syncwarp();
                    if (p.dbg) eS += clock64() - ec0;
                }
            };
            const uint32_t uni = (meta[10] >> (8 * part)) & 0xFFu;  // TileMeta::uni
            // vector stores: the whole store group uniform with the same shape (16 warps: all four parts of the tile),
            // a compiled output set (not the run-time variant), 16-byte aligned row pieces in every array
            const bool vec = OUT != 0 && p.vecStores && uni != 0 && (NE == 8 || meta[10] == uni * 0x01010101u) &&
                             (p.ld % 4) == 0 && (mixBase % 4) == 0 && (reinterpret_cast<uintptr_t>(gBase) % 16) == 0 &&
                             (bestDelta % 16) == 0 && (scoreDelta % 16) == 0;
            auto run = [&](auto spTag) {
                if constexpr (OUT != 0 && decltype(spTag)::value > 0 && (SEGS / (decltype(spTag)::value > 0 ? decltype(spTag)::value : 1)) % 4 == 0) {
                    if (vec)
                        return tiles(spTag, std::true_type{});
                }
                tiles(spTag, std::false_type{});
            };
            switch (uni) {
                case 1: run(std::integral_constant<int, 1>{}); break;
                case 2: run(std::integral_constant<int, 2>{}); break;
                case 4: run(std::integral_constant<int, 4>{}); break;
                case 8: run(std::integral_constant<int, (SEGS >= 8 ? 8 : 0)>{}); break;
                default: run(std::integral_constant<int, 0>{}); break;
            }
            t += f1 - f0;
            if (lane == 0)
                mbar_arrive(smem_u32(&emptyB[b]));
        }
        if (p.dbg && lane == 0) {  // per epilogue warp: waits, phases, whole loop
            unsigned long long* d = p.dbg + blockIdx.x * 128 + 16 + (warp - kEpiWarp0) * 6;
            d[0] = eW, d[1] = eL, d[2] = eB, d[3] = eS, d[4] = clock64() - eStart, d[5] = eI;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmemBase));
    }
}

}  // namespace

// ---- host side ----------------------------------------------------------------------------------------
static uint32_t qSmemNeed(uint32_t bTileBytes, uint32_t aTileBytes, uint32_t buffersB, uint32_t stagesA,
                          uint32_t pitch, uint32_t epiWarps) {
    const uint32_t bBytes = (bTileBytes + 1023u) & ~1023u;
    // 8 staging blocks per plane either way: 8 warps x 1, or 4 lane quarters x 2; eight-warp epilogues come in two groups
    return 1024 + buffersB * bBytes + stagesA * aTileBytes + (epiWarps == 8 ? 2 : 1) * 2 * 8 * 32 * pitch * 4 + 512;
}

// Packs the quantized model for the tensor kernel; leaves h->qtensor.valid false when the model does not meet the
// requirements in the header comment (the dp4a kernel then serves it).  c2[g] = qconst[g] + |qmean_g|^2 (wrapping).
int gmm_qtc_pack(gmm_handle* h, const uint32_t* offsets, const uint8_t* qmean, const int32_t* c2) {
    QTensorModel&     tm = h->qtensor;
    const QuantModel& qm = h->quant;
    tm                   = QTensorModel();
    const uint32_t M = qm.M, G = qm.G, Dpad = qm.Dpad;
    const char*    env = getenv("GMM_QUANT_KERNEL");
    if (G == 0 || M == 0 || qm.C != 1 || (env && !strcmp(env, "dp4a")))
        return GMM_OK;
    // K = Dpad + S bytes with S = 16 or 32 spare bytes for the constants' digits; the B image holds one more k-step
    uint32_t K = (Dpad + 31) / 32 * 32;
    if (K - Dpad < 16)
        K += 32;
    if (K > 128)
        return GMM_OK;  // qtc_prep_kernel: eight 16-byte chunks per frame
    const uint32_t S = K - Dpad;
    tm.K          = K;
    tm.kSteps     = K / 32;
    tm.nAtoms     = (tm.kSteps * 32 + kQSwz - 1) / kQSwz;
    tm.nAtomsB    = ((tm.kSteps + 1) * 32 + kQSwz - 1) / kQSwz;
    tm.aTileBytes = tm.nAtoms * kTileM * kQSwz;
    tm.tileBytes  = tm.nAtomsB * kTileN * kQSwz + kQMetaBytes;
    // value ranges: the constants must fit the spare bytes, the key 32 bits, and nothing may come near wrap-around
    const int64_t dmax = (int64_t)Dpad * 255 * 255;
    int64_t       cmin = INT64_MAX, cmax = INT64_MIN;
    for (uint32_t g = 0; g < G; ++g) {
        cmin = std::min<int64_t>(cmin, c2[g]);
        cmax = std::max<int64_t>(cmax, c2[g]);
    }
    const int64_t cap = (int64_t)255 * 255 * (S - 1) + 254;  // one set of spare bytes: 255 * (b_0 + .. + b_{S-2}) + b_{S-1}
    if (cmin < -(1ll << 30) || cmax > (1ll << 30) || cmax - cmin > 2 * cap || 2 * dmax + (cmax - cmin) >= (1ll << 24))
        return GMM_OK;

    // ---- column assignment: densities of a mixture consecutive, padded to 8, never straddling a 128-column half
    struct Col {
        uint32_t g;  // density, or 0xFFFFFFFF for a pad / filler column
        uint32_t slot;
    };
    uint32_t maxN = 0;
    for (uint32_t m = 0; m < M; ++m)
        maxN = std::max(maxN, offsets[m + 1] - offsets[m]);
    const char* envW = getenv("GMM_QTC_EPI_WARPS");
    // 8 epilogue warps (128-column parts) by default: the 16-warp variant is faster on small batches (48 vs 41 M
    // frames/s at 8000 frames, 4501x16) but its 16-byte output pieces collapse on large ones (20 vs 38 M at 64000)
    // default: 128-column parts = eight-warp epilogue groups, two of them on alternate tiles (GMM_QTC_GROUPS=1: one);
    // GMM_QTC_EPI_WARPS=16 at upload: 64-column parts, one group of sixteen warps (needs mixtures of <= 64 columns)
    tm.partW         = (maxN <= 64 && envW && atoi(envW) == 16) ? 64 : 128;
    const uint32_t        partW = tm.partW, partsPerTile = kTileN / partW;
    std::vector<Col>      cols;
    std::vector<TileMeta> metas;
    std::vector<uint32_t> mixPerPart;
    auto ensurePart = [&](uint32_t pt) {
        while (metas.size() * partsPerTile <= pt) {
            TileMeta z = {0, 0, {0, 0, 0, 0}, {0, 0, 0, 0}, 0};
            metas.push_back(z);
        }
        while (mixPerPart.size() <= pt)
            mixPerPart.push_back(0);
    };
    const Col pad = {0xFFFFFFFFu, 0};
    for (uint32_t m = 0; m < M; ++m) {
        const uint32_t n      = offsets[m + 1] - offsets[m];
        const uint32_t padded = std::max<uint32_t>(8, (n + 7) / 8 * 8);
        if (padded > partW)
            return GMM_OK;  // slot would not fit 8 bits of the key / the column part
        const uint32_t inPart = (uint32_t)(cols.size() % partW);
        if (inPart + padded > partW)
            cols.insert(cols.end(), partW - inPart, pad);
        const uint32_t pt = (uint32_t)(cols.size() / partW);
        ensurePart(pt);
        TileMeta& mt = metas[pt / partsPerTile];
        if (mixPerPart[pt] == 0)
            mt.mixBase[pt % partsPerTile] = m;
        mixPerPart[pt]++;
        const uint32_t seg0 = (uint32_t)((cols.size() % kTileN) / 8);
        for (uint32_t i = 0; i < padded; ++i) {
            Col c = pad;
            if (i < n) {
                c.g    = offsets[m] + i;
                c.slot = i;
            }
            cols.push_back(c);
        }
        mt.segEnd |= 1u << (seg0 + padded / 8 - 1);
        if (n == 0)
            mt.segEmpty |= 1u << (seg0 + padded / 8 - 1);
    }
    while (cols.size() % kTileN)
        cols.push_back(pad);
    tm.nTiles = (uint32_t)(cols.size() / kTileN);
    ensurePart(tm.nTiles * partsPerTile - 1);
    for (TileMeta& mt : metas) {
        const uint32_t segsPerPart = partW / 8;
        for (uint32_t pp = 0; pp < partsPerTile; ++pp) {
            const uint32_t mask = (1u << segsPerPart) - 1u;
            const uint32_t ends = (mt.segEnd >> (pp * segsPerPart)) & mask, empt = (mt.segEmpty >> (pp * segsPerPart)) & mask;
            for (uint32_t SP = 1; SP <= 8 && SP <= segsPerPart; SP *= 2) {
                uint32_t pat = 0;
                for (uint32_t b = SP - 1; b < segsPerPart; b += SP)
                    pat |= 1u << b;
                if (ends == pat && empt == 0)
                    mt.uni |= SP << (8 * pp);
            }
        }
    }
    tm.maxMixPerPart = 1;  // mixtures per store group: a part (8 warps), a whole tile (16 warps)
    for (size_t i = 0; i < mixPerPart.size(); i += (partW == 64 ? partsPerTile : 1)) {
        uint32_t v = 0;
        for (size_t j = i; j < mixPerPart.size() && j < i + (partW == 64 ? partsPerTile : 1); ++j)
            v += mixPerPart[j];
        tm.maxMixPerPart = std::max(tm.maxMixPerPart, v);
    }

    // ---- shared-memory configuration ----
    const uint32_t pitch = (tm.maxMixPerPart + 3) / 4 * 4 + 5, epiWarps = 4 * partsPerTile;
    for (uint32_t nb = 2; nb >= 1 && !tm.buffersB; --nb)
        for (uint32_t sa = kMaxStagesA; sa >= 2; --sa)
            if (qSmemNeed(tm.tileBytes, tm.aTileBytes, nb, sa, pitch, epiWarps) <= kSmemLimit) {
                tm.buffersB = nb;
                tm.stagesA  = sa;
                break;
            }
    if (!tm.buffersB)
        return GMM_OK;

    // ---- tile images ----
    std::vector<uint8_t> img((size_t)tm.nTiles * tm.tileBytes, 0);
    auto at = [&](uint8_t* base, uint32_t r, uint32_t byte) -> uint8_t& {  // byte `byte` of operand row r (64-byte atoms)
        return base[(size_t)(byte / kQSwz) * kTileN * kQSwz + swz_offset(r, (byte % kQSwz) / 16, kQSwz) + (byte % 16)];
    };
    auto digits = [&](uint8_t* base, uint32_t r, uint32_t byte0, int64_t v) {  // v = 255 * (b_0 + .. + b_{S-2}) + b_{S-1}
        int64_t hi = v / 255;
        for (uint32_t i = 0; i + 1 < S; ++i) {
            const int64_t b = std::min<int64_t>(255, hi);
            at(base, r, byte0 + i) = (uint8_t)b;
            hi -= b;
        }
        at(base, r, byte0 + S - 1) = (uint8_t)(v % 255);
    };
    const uint32_t last = (tm.kSteps - 1) * 32, var = tm.kSteps * 32;  // byte offsets of the last k-step and its variant
    for (uint32_t t = 0; t < tm.nTiles; ++t) {
        uint8_t*  base = img.data() + (size_t)t * tm.tileBytes;
        uint32_t* kc   = reinterpret_cast<uint32_t*>(base + (size_t)tm.nAtomsB * kTileN * kQSwz);
        TileMeta* mOut = reinterpret_cast<TileMeta*>(reinterpret_cast<uint8_t*>(kc) + kTileN * 4);
        *mOut          = metas[t];
        for (uint32_t r = 0; r < (uint32_t)kTileN; ++r) {
            const Col& col = cols[(size_t)t * kTileN + r];
            if (col.g == 0xFFFFFFFFu) {
                kc[r] = kNoKey;  // zero operand row: acc = 0, key 0
                continue;
            }
            kc[r] = 255u - col.slot;
            const uint8_t* mu = qmean + (size_t)col.g * Dpad;
            for (uint32_t d = 0; d < Dpad; ++d) {
                at(base, r, d) = mu[d];
                if (d >= last)
                    at(base, r, var + (d - last)) = mu[d];  // the variant repeats the last k-step's data bytes
            }
            const int64_t E = cmax - (int64_t)c2[col.g], E1 = std::min(E, cap);
            digits(base, r, Dpad, E1);
            digits(base, r, var + (Dpad - last), E - E1);
        }
    }
    tm.qBase = (uint32_t)cmax;
    // |q| <= max(|cmin - Dmax|, |cmax + Dmax|): q = qconst + |qmean - qx|^2 with the distance in [0, Dmax] and
    // qconst = c2 - |qmean|^2 >= cmin - Dmax
    tm.qAbsMax = (uint32_t)std::min<int64_t>(0x7FFFFFFF, std::max<int64_t>(std::llabs(cmin - dmax), std::llabs(cmax + dmax)));
    int rc;
    if ((rc = h->upload((void**)&tm.bTiles, img.data(), img.size(), true)))
        return rc;
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    tm.valid = true;
    return GMM_OK;
}

static int qtc_launch(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best, float* d_score,
                      uint8_t* d_qx, size_t ld);

int gmm_qtc_score(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best, float* d_score,
                  uint8_t* d_qx, size_t ld) {
    // launches of at most kChunk frames, as in gmm_tc_score (49.6 M frames/s at 64000 frames per launch, 32.5 M in
    // one 512000-frame launch on the 4501x16 model)
    static const uint32_t kChunk = [] {
        const char* e = getenv("GMM_QTC_CHUNK");
        const long  v = e ? atol(e) : 65536;
        return (uint32_t)std::max<long>(kTileM, v / kTileM * kTileM);
    }();
    const QuantModel& qm = h->quant;
    for (uint32_t t0 = 0; t0 < T; t0 += kChunk) {
        const size_t o  = (size_t)t0 * ld;
        const int    rc = qtc_launch(h, std::min(kChunk, T - t0), d_x + (size_t)t0 * qm.D, d_q ? d_q + o : nullptr,
                                     d_best ? d_best + o : nullptr, d_score ? d_score + o : nullptr,
                                     d_qx ? d_qx + (size_t)t0 * qm.C * qm.Dpad : nullptr, ld);
        if (rc)
            return rc;
    }
    return GMM_OK;
}

static int qtc_launch(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best, float* d_score,
                      uint8_t* d_qx, size_t ld) {
    const QTensorModel& tm = h->qtensor;
    const QuantModel&   qm = h->quant;
    if (T == 0)
        return GMM_OK;
    const uint32_t FT = (T + kTileM - 1) / kTileM;
    void *         aT = nullptr, *rT = nullptr;
    int            rc;
    if ((rc = h->reserve(SCR_A, (size_t)FT * tm.aTileBytes, &aT)))
        return rc;
    if ((rc = h->reserve(SCR_ROW, (size_t)FT * kTileM * 4, &rT)))
        return rc;
    {
        const int blocks = (int)std::min<uint32_t>((FT * kTileM + 31) / 32, (uint32_t)h->smCount * 16);
        qtc_prep_kernel<<<blocks, 256, 0, h->stream>>>(d_x, T, qm.D, qm.Dpad, tm.K, FT, qm.isrS, tm.nAtoms, (uint8_t*)aT,
                                                       (uint32_t*)rT, d_qx);
        h->launches++;
        GMM_CUDA_TRY(h, cudaGetLastError());
    }
    QKParams p;
    p.bTiles         = tm.bTiles;
    p.aTiles         = (const uint8_t*)aT;
    p.rowXX          = (const uint32_t*)rT;
    p.q              = d_q;
    p.best           = d_best;
    p.score          = d_score;
    p.ld             = ld;
    p.scalingSquared = qm.finalScoreArg();
    p.T              = T;
    p.nTiles         = tm.nTiles;
    p.frameTiles     = FT;
    p.kSteps         = tm.kSteps;
    p.atoms          = tm.nAtoms;
    p.atomsB         = tm.nAtomsB;
    p.aTileBytes     = tm.aTileBytes;
    p.bTileBytes     = tm.tileBytes;
    p.stagesA        = tm.stagesA;
    p.buffersB       = tm.buffersB;
    p.stagePitch     = (tm.maxMixPerPart + 3) / 4 * 4 + 5;  // odd, and >= W + 4 of the vector store path
    p.qBase          = tm.qBase;
    p.dbg            = nullptr;
    {
        const char* e = getenv("GMM_QTC_STORES");
        p.vecStores   = (e && !strcmp(e, "scalar")) ? 0u : 1u;
    }
    const bool debug = getenv("GMM_QTC_DEBUG") != nullptr;
    if (debug) {
        void* d = nullptr;
        if ((rc = h->reserve(SCR_MISC, (size_t)h->smCount * 1024, &d)))
            return rc;
        GMM_CUDA_TRY(h, cudaMemsetAsync(d, 0, (size_t)h->smCount * 1024, h->stream));
        p.dbg = (unsigned long long*)d;
    }
    const uint32_t epiWarps = 4 * (kTileN / tm.partW);
    const size_t   smemBytes = qSmemNeed(tm.tileBytes, tm.aTileBytes, tm.buffersB, tm.stagesA, p.stagePitch, epiWarps);
    // frame range split so that the items fill the SMs evenly (round-robin items keep adjacent B tiles together in
    // time, see gmm_tc.cu)
    const double bCost = tm.buffersB == 2 ? 0.5 : 2.0;
    uint32_t     bestS = 1;
    double       bestCost = 1e300;
    for (uint32_t S = 1; S <= std::min<uint32_t>(FT, 64); ++S) {
        const uint32_t per = (FT + S - 1) / S;
        if ((uint64_t)per * (S - 1) >= FT)
            continue;
        const uint64_t items  = (uint64_t)tm.nTiles * S;
        const uint64_t rounds = (items + h->smCount - 1) / h->smCount;
        const double   cost   = (double)rounds * (per + bCost);
        if (cost < bestCost - 1e-9) {
            bestCost = cost;
            bestS    = S;
        }
    }
    if (const char* e = getenv("GMM_QTC_SPLIT"))  // experiments: frame ranges per B tile
        bestS = std::max(1u, std::min<uint32_t>(FT, (uint32_t)atoi(e)));
    p.split        = bestS;
    const int grid = (int)std::min<uint64_t>((uint64_t)h->smCount, (uint64_t)tm.nTiles * bestS);
    // The last B tile is usually a partial one (pad columns: the ragged epilogue path, 1.2x the time of the others on the
    // 4501x16 model, where it made CTA 133 finish 20 us after everybody else at 8000 frames).  With whole frame ranges per
    // item it trades places with a first-round tile of a CTA that has one item fewer.
    p.swapTile = tm.nTiles - 1;
    if (bestS == 1 && tm.nTiles > (uint32_t)grid && tm.nTiles % grid != 0)
        p.swapTile = grid - 1;
    // Output set fixed at compile time for the combinations callers use (tests / bench: q + arg-min; the RASR plugin:
    // score, or score + arg-min; everything).  The final score there is ONE f32 division q / (2 s^2): for |q| < 2^24 the
    // integer is exact in f32 and the correctly rounded f32 quotient equals the reference's f32(0.5 * q / s^2) evaluated
    // in f64 (SimdFeatureScorer.cc:135-145) -- q / d with a 24-bit d lies at least 2^-49 (relative) away from every f32
    // rounding boundary it does not hit exactly, so the intermediate f64 rounding (2^-53) cannot change the result.
    // tm.qAbsMax bounds |q| over all frames (packing); larger constants keep the f64 expression (OUT = 0).
    int out = (d_q ? 1 : 0) | (d_best ? 2 : 0) | (d_score ? 4 : 0);
    p.twoScalingSquared = 2.0f * qm.scalingSquared;
    const bool generic = getenv("GMM_QTC_GENERIC") != nullptr;  // tests: force the run-time variant
    if (generic || (out != 3 && out != 4 && out != 6 && out != 7) ||
        ((out & 4) && !qm.scoreRule &&
         (tm.qAbsMax >= (1u << 24) || !std::isfinite(p.twoScalingSquared) || !(qm.scalingSquared > 0.f))))
        out = 0;
    void (*kernel)(const QKParams) = nullptr;
    uint32_t groups = 1;
    {
        const char* e = getenv("GMM_QTC_GROUPS");
        if (epiWarps == 8 && !(e && atoi(e) == 1))
            groups = 2;
    }
    p.keyMul = 256;
    p.dbgSkip = getenv("GMM_QTC_DEBUG_SKIP") ? (uint32_t)atoi(getenv("GMM_QTC_DEBUG_SKIP")) : 0u;
    if (epiWarps == 16)
        kernel = out == 3 ? qtc_score_kernel<16, 1, 3> : out == 4 ? qtc_score_kernel<16, 1, 4> : out == 6 ? qtc_score_kernel<16, 1, 6>
               : out == 7 ? qtc_score_kernel<16, 1, 7> : qtc_score_kernel<16, 1, 0>;
    else if (groups == 2)
        kernel = out == 3 ? qtc_score_kernel<8, 2, 3> : out == 4 ? qtc_score_kernel<8, 2, 4> : out == 6 ? qtc_score_kernel<8, 2, 6>
               : out == 7 ? qtc_score_kernel<8, 2, 7> : qtc_score_kernel<8, 2, 0>;
    else
        kernel = out == 3 ? qtc_score_kernel<8, 1, 3> : out == 4 ? qtc_score_kernel<8, 1, 4> : out == 6 ? qtc_score_kernel<8, 1, 6>
               : out == 7 ? qtc_score_kernel<8, 1, 7> : qtc_score_kernel<8, 1, 0>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    GMM_CUDA_TRY(h, e);
    h->timeBegin();
    kernel<<<grid, 32 * (kEpiWarp0 + epiWarps * groups), smemBytes, h->stream>>>(p);
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    if (debug) {  // development aid: where the MMA thread and epilogue warp 0 spend their cycles
        std::vector<unsigned long long> d((size_t)h->smCount * 128);
        GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
        GMM_CUDA_TRY(h, cudaMemcpy(d.data(), p.dbg, d.size() * 8, cudaMemcpyDeviceToHost));
        double tot = 0, mW = 0, mI = 0, tiles = 0, ns = 0;
        unsigned long long g0 = ~0ull, g1 = 0;
        for (int c = 0; c < grid; ++c) {
            tot += d[c * 128], mW += d[c * 128 + 1], mI += d[c * 128 + 2], tiles += d[c * 128 + 4];
            ns += (double)(d[c * 128 + 6] - d[c * 128 + 5]);
            g0 = std::min(g0, d[c * 128 + 5]), g1 = std::max(g1, d[c * 128 + 6]);
        }
        fprintf(stderr, "[gmm_qtc] MMA loops: %.1f us first start to last end, SM clock %.0f MHz (cycles / globaltimer)\n",
                (double)(g1 - g0) * 1e-3, ns > 0 ? tot / ns * 1e3 : 0.0);
        {  // per CTA: start offset, duration, tiles -- by number of tiles
            double dmin[2] = {1e30, 1e30}, dmax[2] = {0, 0}, dsum[2] = {0, 0}, smax = 0, emin = 1e30;
            int    n[2] = {0, 0};
            unsigned long long tmax = 0;
            for (int c = 0; c < grid; ++c)
                tmax = std::max(tmax, d[c * 128 + 4]);
            for (int c = 0; c < grid; ++c) {
                const int    k   = d[c * 128 + 4] == tmax ? 1 : 0;
                const double dur = (double)(d[c * 128 + 6] - d[c * 128 + 5]) * 1e-3;
                dmin[k] = std::min(dmin[k], dur), dmax[k] = std::max(dmax[k], dur), dsum[k] += dur, n[k]++;
                smax = std::max(smax, (double)(d[c * 128 + 5] - g0) * 1e-3);
                emin = std::min(emin, (double)(d[c * 128 + 6] - g0) * 1e-3);
            }
            std::vector<int> order(grid);
            for (int c = 0; c < grid; ++c)
                order[c] = c;
            std::sort(order.begin(), order.end(), [&](int a, int b) {
                return d[a * 128 + 6] - d[a * 128 + 5] > d[b * 128 + 6] - d[b * 128 + 5];
            });
            for (int i = 0; i < std::min(grid, 6); ++i) {
                const int c = order[i];
                fprintf(stderr, "[gmm_qtc]   slowest %d: CTA %d on SM %llu, %llu tiles, %.1f us, wait-ready %.0f issue %.0f per tile\n", i, c,
                        d[c * 128 + 7], d[c * 128 + 4], (double)(d[c * 128 + 6] - d[c * 128 + 5]) * 1e-3,
                        (double)d[c * 128 + 1] / std::max<double>(1, d[c * 128 + 4]), (double)d[c * 128 + 2] / std::max<double>(1, d[c * 128 + 4]));
            }
            fprintf(stderr, "[gmm_qtc]   %d CTAs with %llu tiles: loop %.1f / %.1f / %.1f us (min / mean / max); %d CTAs with fewer: "
                    "%.1f / %.1f / %.1f us; latest start +%.1f us, earliest end +%.1f us\n",
                    n[1], tmax, dmin[1], dsum[1] / std::max(1, n[1]), dmax[1], n[0], n[0] ? dmin[0] : 0.0,
                    dsum[0] / std::max(1, n[0]), dmax[0], smax, emin);
        }
        tiles = std::max(1.0, tiles);
        fprintf(stderr,
                "[gmm_qtc] grid %d split %u stagesA %u buffersB %u kSteps %u epiWarps %u x %u out %d | cycles per tile: MMA "
                "thread total %.0f = wait-ready %.0f + issue %.0f\n",
                grid, p.split, p.stagesA, p.buffersB, p.kSteps, epiWarps, groups, out, tot / tiles, mW / tiles, mI / tiles);
        for (uint32_t w = 0; w < epiWarps * groups; ++w) {  // per CTA tile (a group of two sees every other tile)
            double v[6] = {0, 0, 0, 0, 0, 0};
            for (int c = 0; c < grid; ++c)
                for (int k = 0; k < 6; ++k)
                    v[k] += d[c * 128 + 16 + w * 6 + k];
            fprintf(stderr,
                    "[gmm_qtc]   epilogue warp %2u: wait-for-MMA %.0f + loads and reduction %.0f + barrier %.0f + stores %.0f "
                    "= %.0f of loop %.0f (B-tile wait %.0f)\n",
                    w, v[0] / tiles, v[1] / tiles, v[2] / tiles, v[3] / tiles, (v[0] + v[1] + v[2] + v[3]) / tiles,
                    v[4] / tiles, v[5] / tiles);
        }
    }
    return GMM_OK;
}
