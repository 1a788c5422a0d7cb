// Tensor-core float scorer for sm_100a: the diagonal-Gaussian distance as one dense contraction on tcgen05
// with accumulators in TMEM and a fused per-mixture min / log-sum-exp epilogue, so density-level scores
// never reach HBM.  Replaces the inner loops of the reference's float scorers
// (/root/reference/src/Mm/GaussDiagonalMaximumFeatureScorer.cc:116-218 max, :239-289 sum).
//
// Math (pooled covariance, C == 1; "whitened" form):
//     a_t = (x_t - ctr) * isr            b_g = (mu_g - ctr) * isr
//     w2_g + L + dist(x_t, g) = |a_t|^2 + ( c_g - 2 a_t.b_g ),   c_g = |b_g|^2 + L + w2_g
//   general covariance (C > 1):  A row = [ (x s)^2 , x s ],  B row = [ iv/s^2 , -2 mu iv / s ],
//     c_g = sum mu^2 iv + L_c + w2_g,  row term 0  (s = per-dimension scale that keeps fp16 in range).
// Split precision: every operand value v is stored as fp16 hi = rn(v), lo = rn(v - hi) (22 significant bits);
// the contraction runs hi.hi + lo.hi + hi.lo with f32 accumulation in TMEM (kind::f16, 2x the TF32 rate).
// The per-Gaussian constant c_g rides in three spare K columns as fp16 pieces c0+c1+c2 (33 bits) against ones
// on the frame side, so the accumulator already holds c_g - 2 a.b and the max-approx epilogue is a bare min
// tree.  Zero-weight densities (c = +inf in the reference) are dropped at packing time: they never win a min.
//
// Layout.  Both operands live in HBM as ready-made shared-memory tile images (K-major, 128B- or 64B-swizzled
// canonical UMMA layout), so one cp.async.bulk (TMA unit, SASS UBLKCP) per tile stages them:
//   B tile  = 256 packed columns (Gaussians):  [atoms][256 rows][SWZ bytes] | slot[256] u16 | TileMeta
//   A tile  = 128 frames:                       [atoms][128 rows][SWZ bytes]
// Packed columns: the live densities of a mixture are consecutive, padded to a multiple of 8 columns (a pad
// repeats the mixture's first live column); a mixture never straddles a 128-column half tile.  TileMeta: bit s
// of segEnd says "8-column segment s closes a mixture", plus live-column counts and the first mixture per half.
//
// Kernel: persistent, one CTA per SM, 12 warps: warp 0 bulk-copy producer, warp 1 MMA issuer (one elected
// thread, M=128 N=256 K=16 per instruction), warp 2 TMEM allocator, warps 4-11 epilogue (TMEM lane quarter =
// warp % 4, column half = (warp-4)/4).  Work item = (B tile, range of frame tiles); B stays in shared memory for
// the whole item while A tiles stream through a ring; two 256-column accumulators in TMEM double-buffer MMA
// against the epilogue.
#include <cuda_fp16.h>
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "gmm_internal.cuh"

namespace {

constexpr int kTileN       = 256;  // packed columns per B tile (UMMA N)
constexpr int kTileM       = 128;  // frames per A tile (UMMA M, TMEM lanes)
constexpr int kThreads     = 384;
constexpr int kEpiWarp0    = 4;
constexpr int kMaxStagesA  = 4;
constexpr int kMetaBytes   = kTileN * 2 + 32;  // slot[256] u16, TileMeta
constexpr uint32_t kSmemLimit = 232448;                     // 227 KB

struct TileMeta {
    uint32_t segEnd;       // bit s: 8-column segment s closes a mixture
    uint32_t segEmpty;     // bit s (on a closing segment): the mixture has no live density at all
    uint32_t mixBase0, mixBase1;  // first mixture of column half 0 / 1
    uint32_t live[4];      // 4 bits per segment: live columns - 1 (pads repeat the first live column)
};

// ---- PTX wrappers ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "WAIT_%=:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
            "@p bra DONE_%=;\n\t"
            "bra WAIT_%=;\n\t"
            "DONE_%=:\n\t}" ::"r"(bar),
            "r"(parity)
            : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(dTmem),
            "l"(aDesc), "l"(bDesc), "r"(idesc), "r"(accumulate)
            : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
    asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
              "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
              "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr)
            : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// UMMA shared-memory descriptor, K-major, swizzled canonical layout (8-row groups of SWZ-byte rows).
template<int SWZ>
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    constexpr uint64_t layout = SWZ == 128 ? 2ull : 4ull;  // SWIZZLE_128B : SWIZZLE_64B
    constexpr uint64_t sbo    = (8 * SWZ) >> 4;            // stride between 8-row groups
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | (sbo << 32) | (1ull << 46) | (layout << 61);
}

struct KParams {
    const uint8_t* bTiles;
    const uint8_t* aTiles;
    const float2*  rowInfo;  // per frame row: {row term |a|^2, power-of-two row scale}
    float*         scores;
    uint32_t*      best;
    size_t         ld;
    uint32_t       T;
    uint32_t       nTiles, frameTiles, split;  // items = nTiles * split
    uint32_t       kSteps;                     // k-steps (16 elements) per precision part
    uint32_t       atoms;                      // swizzle atoms per operand row
    uint32_t       aTileBytes, bTileBytes;     // bTileBytes includes c / slot / meta
    uint32_t       stagesA, buffersB;
    uint32_t       stagePitch;                 // floats per staging row (maxMixPerHalf + 1)
};

struct EpiState {
    float    run, sum;
    uint32_t idx, cnt;
};

// One 32-column chunk (four 8-column segments) of a row's accumulator.  The accumulator already holds
// u / rowScale with u = c - 2 a.b (the constant rides in spare K columns of the contraction), so the maximum
// approximation is a bare min tree over TMEM values; a closing segment stages the mixture's score.
template<int MODE, bool BEST>
__device__ __forceinline__ void epi_chunk(const uint32_t* v, int ch, int half, const uint16_t* __restrict__ slots,
                                          uint32_t segEnd, uint32_t segEmpty, uint64_t liveBits, float rt, float rs,
                                          float* stg, uint32_t* stgB, EpiState& st) {
    constexpr float kLn2 = 0.69314718055994530942f;
    const float     hl2e = 0.72134752044448170368f * rs;  // 0.5 * log2(e) * rowScale
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const int seg = ch * 4 + s;
        float     u[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
            u[j] = __uint_as_float(v[s * 8 + j]);
        if (MODE == 0 && !BEST) {
            float m8 = fminf(fminf(fminf(u[0], u[1]), fminf(u[2], u[3])), fminf(fminf(u[4], u[5]), fminf(u[6], u[7])));
            st.run   = fminf(st.run, m8);
        }
        else if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (u[j] < st.run) {  // pads repeat the first live column: never strictly smaller
                    st.run = u[j];
                    st.idx = (uint32_t)(half * 128 + seg * 8 + j);
                }
        }
        else {
            const int live = (int)((liveBits >> (seg * 4)) & 15ull) + 1;  // uniform
            float     m8 = u[0];
            int       j8 = 0;
#pragma unroll
            for (int j = 1; j < 8; ++j)
                if (u[j] < m8) {
                    m8 = u[j];
                    j8 = j;
                }
            // exp(-0.5 * 40) < 2^-28: a segment that far above the running best of every frame of the warp
            // cannot change an f32 sum that is >= 1
            if (__any_sync(0xffffffffu, !((m8 - st.run) * rs >= 40.f))) {
                if (m8 < st.run) {
                    st.sum *= ex2((m8 - st.run) * hl2e);  // run == inf -> factor 0 (sum is 0 anyway)
                    st.run = m8;
                    st.idx = (uint32_t)(half * 128 + seg * 8 + j8);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (j < live)
                        st.sum += ex2((st.run - u[j]) * hl2e);
            }
        }
        if ((segEnd >> seg) & 1u) {  // uniform across the warp
            const bool  empty = (segEmpty >> seg) & 1u;
            const float best  = empty ? FLT_MAX : fminf(st.run * rs, FLT_MAX);
            float       sc    = 0.5f * (best + rt);
            if (MODE == 1)
                sc = empty ? INFINITY : sc - kLn2 * __log2f(st.sum);
            stg[st.cnt] = sc;
            if (BEST)
                stgB[st.cnt] = (empty || st.idx == 0xFFFFFFFFu) ? 0xFFFFFFFFu : (uint32_t)slots[st.idx];
            ++st.cnt;
            st.run = INFINITY;
            st.sum = 0.f;
            st.idx = 0xFFFFFFFFu;
        }
    }
}

// MODE 0 = max, 1 = sum.  BEST: also write the index of the winning density.
template<int SWZ, int MODE, bool BEST>
__global__ void __launch_bounds__(kThreads, 1) tc_score_kernel(const KParams p) {
    extern __shared__ uint8_t smemRaw[];
    // swizzled operand tiles need 1024-byte alignment of the shared-memory address itself
    uint8_t*  smem = smemRaw + ((1024u - (smem_u32(smemRaw) & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // carve shared memory: [B buffers][A stages][staging][barriers]
    const uint32_t bBytes    = (p.bTileBytes + 1023u) & ~1023u;
    uint8_t*       sB        = smem;
    uint8_t*       sA        = sB + (size_t)p.buffersB * bBytes;
    float*         sStage    = reinterpret_cast<float*>(sA + (size_t)p.stagesA * p.aTileBytes);
    const uint32_t stageWarp = 32 * p.stagePitch;  // floats per warp
    uint32_t*      sStageB   = reinterpret_cast<uint32_t*>(sStage + 8 * stageWarp);
    uint64_t*      bars      = reinterpret_cast<uint64_t*>(sStageB + (BEST ? 8 * stageWarp : 0));
    // barrier indices
    uint64_t* fullA     = bars;                    // [kMaxStagesA]
    uint64_t* emptyA    = bars + kMaxStagesA;      // [kMaxStagesA]
    uint64_t* fullB     = bars + 2 * kMaxStagesA;  // [2]
    uint64_t* emptyB    = fullB + 2;               // [2]
    uint64_t* tmemFull  = emptyB + 2;              // [2]
    uint64_t* tmemEmpty = tmemFull + 2;            // [2]
    uint32_t* tmemSlot  = reinterpret_cast<uint32_t*>(tmemEmpty + 2);

    if (threadIdx.x == 0) {
        for (int i = 0; i < kMaxStagesA; ++i) {
            mbar_init(smem_u32(&fullA[i]), 1);
            mbar_init(smem_u32(&emptyA[i]), 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(smem_u32(&fullB[i]), 1);
            mbar_init(smem_u32(&emptyB[i]), 1 + 8);  // MMA commit + 8 epilogue warps
            mbar_init(smem_u32(&tmemFull[i]), 1);
            mbar_init(smem_u32(&tmemEmpty[i]), 8);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
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
        // ===== producer: one lane issues the bulk copies =====
        if (lane == 0) {
            uint32_t stage = 0, phaseA = 0, itemCount = 0;
            for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
                const uint32_t nt = item / p.split, chunk = item % p.split;
                const uint32_t f0 = chunk * tilesPerItem, f1 = min(p.frameTiles, f0 + tilesPerItem);
                if (f0 >= f1)
                    continue;
                const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
                mbar_wait(smem_u32(&emptyB[b]), (useB & 1) ^ 1);
                mbar_expect_tx(smem_u32(&fullB[b]), p.bTileBytes);
                bulk_g2s(smem_u32(sB + (size_t)b * bBytes), p.bTiles + (size_t)nt * p.bTileBytes, p.bTileBytes,
                         smem_u32(&fullB[b]));
                for (uint32_t f = f0; f < f1; ++f) {
                    mbar_wait(smem_u32(&emptyA[stage]), phaseA ^ 1);
                    mbar_expect_tx(smem_u32(&fullA[stage]), p.aTileBytes);
                    bulk_g2s(smem_u32(sA + (size_t)stage * p.aTileBytes), p.aTiles + (size_t)f * p.aTileBytes,
                             p.aTileBytes, smem_u32(&fullA[stage]));
                    if (++stage == p.stagesA) {
                        stage = 0;
                        phaseA ^= 1;
                    }
                }
            }
        }
    }
    else if (warp == 1) {
        // ===== MMA issuer: one lane issues tcgen05.mma for the whole CTA =====
        if (lane == 0) {
            // instruction descriptor: D=f32, A=B=f16, K-major both, N=256, M=128
            constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(kTileN >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
            constexpr uint32_t kPerAtom = SWZ / 32;  // 16-element k-steps per swizzle atom
            uint32_t stage = 0, phaseA = 0, itemCount = 0, tileCount = 0;
            for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
                const uint32_t chunk = item % p.split;
                const uint32_t f0 = chunk * tilesPerItem, f1 = min(p.frameTiles, f0 + tilesPerItem);
                if (f0 >= f1)
                    continue;
                const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
                mbar_wait(smem_u32(&fullB[b]), useB & 1);
                const uint32_t bAddr = smem_u32(sB + (size_t)b * bBytes);
                for (uint32_t f = f0; f < f1; ++f, ++tileCount) {
                    const uint32_t acc = tileCount & 1, useAcc = tileCount >> 1;
                    mbar_wait(smem_u32(&tmemEmpty[acc]), (useAcc & 1) ^ 1);
                    mbar_wait(smem_u32(&fullA[stage]), phaseA);
                    tc_fence_after();
                    const uint32_t aAddr = smem_u32(sA + (size_t)stage * p.aTileBytes);
                    const uint32_t dTmem = tmemBase + acc * kTileN;
                    uint32_t       first = 1;
                    // hi.hi, lo.hi, hi.lo
#pragma unroll 1
                    for (uint32_t pair = 0; pair < 3; ++pair) {
                        const uint32_t pa = (pair == 1) ? p.kSteps : 0, pb = (pair == 2) ? p.kSteps : 0;
                        for (uint32_t ks = 0; ks < p.kSteps; ++ks) {
                            const uint32_t ka = pa + ks, kb = pb + ks;
                            const uint32_t ao = (ka / kPerAtom) * (kTileM * SWZ) + (ka % kPerAtom) * 32;
                            const uint32_t bo = (kb / kPerAtom) * (kTileN * SWZ) + (kb % kPerAtom) * 32;
                            tc_mma_f16(dTmem, umma_desc<SWZ>(aAddr + ao), umma_desc<SWZ>(bAddr + bo), idesc,
                                       first ? 0u : 1u);
                            first = 0;
                        }
                    }
                    tc_commit(smem_u32(&emptyA[stage]));   // A stage free when these MMAs retire
                    tc_commit(smem_u32(&tmemFull[acc]));   // accumulator ready for the epilogue
                    if (f + 1 == f1)
                        tc_commit(smem_u32(&emptyB[b]));   // B buffer no longer read by the tensor core
                    if (++stage == p.stagesA) {
                        stage = 0;
                        phaseA ^= 1;
                    }
                }
            }
        }
    }
    else if (warp >= kEpiWarp0) {
        // ===== epilogue: TMEM -> registers -> (+c, min / log-sum-exp per mixture) -> staged -> global =====
        const int      e = warp - kEpiWarp0, q = warp & 3, half = e >> 2;
        float*         stg  = sStage + (size_t)e * stageWarp + (size_t)lane * p.stagePitch;
        uint32_t*      stgB = sStageB + (size_t)e * stageWarp + (size_t)lane * p.stagePitch;
        uint32_t       itemCount = 0, tileCount = 0;
        for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
            const uint32_t nt = item / p.split, chunk = item % p.split;
            const uint32_t f0 = chunk * tilesPerItem, f1 = min(p.frameTiles, f0 + tilesPerItem);
            if (f0 >= f1)
                continue;
            (void)nt;
            const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
            mbar_wait(smem_u32(&fullB[b]), useB & 1);
            const uint8_t*  tile  = sB + (size_t)b * bBytes + (size_t)p.atoms * kTileN * SWZ;
            const uint16_t* slots = reinterpret_cast<const uint16_t*>(tile);
            const TileMeta  meta  = *reinterpret_cast<const TileMeta*>(tile + kTileN * 2);
            const uint32_t  segEnd  = (meta.segEnd >> (16 * half)) & 0xFFFFu;
            const uint32_t  mixBase = half ? meta.mixBase1 : meta.mixBase0;
            const uint32_t  segEmpty = (meta.segEmpty >> (16 * half)) & 0xFFFFu;
            const uint64_t  liveBits = half ? ((uint64_t)meta.live[3] << 32 | meta.live[2])
                                            : ((uint64_t)meta.live[1] << 32 | meta.live[0]);
            const uint32_t  nMix    = __popc(segEnd);
            for (uint32_t f = f0; f < f1; ++f, ++tileCount) {
                const uint32_t acc = tileCount & 1, useAcc = tileCount >> 1;
                const uint32_t row = f * kTileM + q * 32 + lane;
                const float2   ri  = p.rowInfo[row];
                const float    rt = ri.x, rs = ri.y;
                mbar_wait(smem_u32(&tmemFull[acc]), useAcc & 1);
                tc_fence_after();
                const uint32_t taddr = tmemBase + ((uint32_t)(q * 32) << 16) + acc * kTileN + half * 128;
                EpiState st;
                st.run = INFINITY, st.sum = 0.f, st.idx = 0xFFFFFFFFu, st.cnt = 0;
                // TMEM -> registers in four 32-column chunks, the next load in flight while one is reduced
                uint32_t va[32], vb[32];
                tmem_ld32(taddr, va);
                tmem_ld_wait();
                tmem_ld32(taddr + 32, vb);
                epi_chunk<MODE, BEST>(va, 0, half, slots, segEnd, segEmpty, liveBits, rt, rs, stg, stgB, st);
                tmem_ld_wait();
                tmem_ld32(taddr + 64, va);
                epi_chunk<MODE, BEST>(vb, 1, half, slots, segEnd, segEmpty, liveBits, rt, rs, stg, stgB, st);
                tmem_ld_wait();
                tmem_ld32(taddr + 96, vb);
                epi_chunk<MODE, BEST>(va, 2, half, slots, segEnd, segEmpty, liveBits, rt, rs, stg, stgB, st);
                tmem_ld_wait();
                // accumulator drained into registers: hand it back to the MMA warp before finishing the tile
                tc_fence_before();
                __syncwarp();
                if (lane == 0)
                    mbar_arrive(smem_u32(&tmemEmpty[acc]));
                epi_chunk<MODE, BEST>(vb, 3, half, slots, segEnd, segEmpty, liveBits, rt, rs, stg, stgB, st);
                __syncwarp();
                // coalesced store of the staged [32 rows x nMix] block
                const uint32_t total = 32 * nMix;
                const float*   wst   = sStage + (size_t)e * stageWarp;
                const uint32_t* wsb  = sStageB + (size_t)e * stageWarp;
                const uint32_t rbase = f * kTileM + q * 32;
                const bool     pow2  = (nMix & (nMix - 1)) == 0;
                const uint32_t sh    = 31 - __clz(nMix | 1);
                for (uint32_t i = lane; i < total; i += 32) {
                    const uint32_t r = pow2 ? (i >> sh) : (i / nMix), j = i - r * nMix;
                    if (rbase + r < p.T) {
                        const size_t o = (size_t)(rbase + r) * p.ld + mixBase + j;
                        p.scores[o]    = wst[r * p.stagePitch + j];
                        if (BEST)
                            p.best[o] = wsb[r * p.stagePitch + j];
                    }
                }
                __syncwarp();
            }
            // this warp no longer reads c / slot / meta of the B buffer
            if (lane == 0)
                mbar_arrive(smem_u32(&emptyB[b]));
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmemBase));
    }
}

// ---- swizzled tile image addressing (host + device) ---------------------------------------------------
// byte offset of 16-byte chunk `c16` of row `row` inside one atom block [rows][SWZ bytes]
__host__ __device__ inline uint32_t swz_offset(uint32_t row, uint32_t c16, uint32_t swz) {
    const uint32_t lin  = row * swz + c16 * 16;
    const uint32_t mask = (swz / 16) - 1;  // 7 for 128B, 3 for 64B
    return lin ^ (((lin >> 7) & mask) << 4);
}

__device__ __forceinline__ void split_fp16(float v, __half& hi, __half& lo) {
    v  = fminf(fmaxf(v, -60000.f), 60000.f);
    hi = __float2half_rn(v);
    lo = __float2half_rn(v - __half2float(hi));
}

// Frame preparation: x [T x D] -> A tile images (fp16 hi | lo parts, swizzled) + per-row {term, scale}.
//   pooled : a = (x - ctr) * isr,                   K = D,   row term = |a|^2
//   general: a = [ ((x - ctr) s)^2 , (x - ctr) s ],  K = 2D,  row term = 0
// A row whose largest |a| would leave the fp16 range is stored divided by a power of two (row scale); the
// epilogue multiplies the per-mixture minimum back, so outlier frames stay exact.  Three spare K columns hold
// 1/rowScale: they multiply the fp16 pieces of the per-Gaussian constant stored in the B operand.
// One warp per frame: lane l owns the 8-element chunk l of each precision part (kSteps <= 16 -> <= 32 chunks).
template<bool POOLED>
__global__ void tc_prep_kernel(const float* __restrict__ x, uint32_t T, uint32_t D, uint32_t frameTiles,
                               const float* __restrict__ ctr, const float* __restrict__ scale, uint32_t kSteps,
                               uint32_t atoms, uint32_t swz, uint8_t* __restrict__ aTiles,
                               float2* __restrict__ rowInfo) {
    const uint32_t chunks = kSteps * 2;  // 8-element chunks per precision part
    const uint32_t epa    = swz / 2;     // fp16 elements per atom row
    const uint32_t lane   = threadIdx.x & 31;
    const uint32_t kel    = POOLED ? D : 2 * D;
    const uint32_t warpsPerGrid = gridDim.x * (blockDim.x >> 5);
    for (uint32_t t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < frameTiles * kTileM; t += warpsPerGrid) {
        float v[8];
        float vmax = 0.f, ssq = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t k = lane * 8 + j;
            v[j]             = 0.f;
            if (t < T && lane < chunks) {
                if (POOLED) {
                    if (k < D) {
                        v[j] = (x[(size_t)t * D + k] - ctr[k]) * scale[k];
                        ssq  = fmaf(v[j], v[j], ssq);
                    }
                }
                else if (k < D) {
                    const float y = (x[(size_t)t * D + k] - ctr[k]) * scale[k];
                    v[j]          = y * y;
                }
                else if (k < 2 * D) {
                    v[j] = (x[(size_t)t * D + (k - D)] - ctr[k - D]) * scale[k - D];
                }
            }
            if (k >= kel && k < kel + 3)
                v[j] = 1.f;  // multiplies the three fp16 pieces of the per-Gaussian constant (stored as 1/rowScale)
            vmax = fmaxf(vmax, fabsf(v[j]));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
            ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
        }
        // power-of-two row scale: largest |a| / rs <= 16384 (non-finite frames are passed through clamped)
        float rs = 1.f;
        if (vmax > 16384.f && vmax < 3.0e38f) {
            int e;
            frexpf(vmax, &e);  // vmax = m * 2^e, m in [0.5, 1)
            rs = ldexpf(1.f, e - 14);
        }
        const float inv = 1.f / rs;
        if (lane == 0)
            rowInfo[t] = make_float2(ssq, rs);
        const uint32_t tile = t / kTileM, row = t % kTileM;
        uint8_t*       base = aTiles + (size_t)tile * atoms * kTileM * swz;
        if (lane < chunks) {
            __align__(16) __half hi[8], lo[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                split_fp16(v[j] * inv, hi[j], lo[j]);
            const uint32_t kHi = lane * 8, kLo = kSteps * 16 + lane * 8;
            *reinterpret_cast<uint4*>(base + (size_t)(kHi / epa) * kTileM * swz +
                                      swz_offset(row, (kHi % epa) / 8, swz)) = *reinterpret_cast<const uint4*>(hi);
            *reinterpret_cast<uint4*>(base + (size_t)(kLo / epa) * kTileM * swz +
                                      swz_offset(row, (kLo % epa) / 8, swz)) = *reinterpret_cast<const uint4*>(lo);
        }
        // zero the unused tail of the last atom
        const uint32_t used = chunks * 2, all = atoms * (swz / 16);  // 16-byte chunks per row
        for (uint32_t c = used + lane; c < all; c += 32)
            *reinterpret_cast<uint4*>(base + (size_t)(c / (swz / 16)) * kTileM * swz +
                                      swz_offset(row, c % (swz / 16), swz)) = make_uint4(0, 0, 0, 0);
    }
}

// host-side fp16 split (round to nearest even), mirrors split_fp16
void host_split(float v, uint16_t& hi, uint16_t& lo) {
    v         = std::min(std::max(v, -60000.f), 60000.f);
    __half h  = __float2half_rn(v);
    float  hf = __half2float(h);
    __half l  = __float2half_rn(v - hf);
    memcpy(&hi, &h, 2);
    memcpy(&lo, &l, 2);
}

template<int SWZ, int MODE, bool BEST>
cudaError_t launch_tc(const KParams& p, int grid, size_t smemBytes, cudaStream_t stream) {
    cudaError_t e = cudaFuncSetAttribute(tc_score_kernel<SWZ, MODE, BEST>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smemBytes);
    if (e != cudaSuccess)
        return e;
    tc_score_kernel<SWZ, MODE, BEST><<<grid, kThreads, smemBytes, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace

// ---- host side: packing --------------------------------------------------------------------------------
static uint32_t smemNeed(uint32_t bTileBytes, uint32_t aTileBytes, uint32_t buffersB, uint32_t stagesA,
                         uint32_t pitch, bool best) {
    const uint32_t bBytes = (bTileBytes + 1023u) & ~1023u;
    return 1024 + buffersB * bBytes + stagesA * aTileBytes + 8 * 32 * pitch * 4 * (best ? 2 : 1) + 512;
}

int gmm_tc_pack(gmm_handle* h, const uint32_t* offsets, const float* mean, const float* isr, const uint32_t* covIdx,
                const float* w2, const float* logNorm, bool force) {
    TensorModel& tm = h->tensor;
    tm              = TensorModel();
    const uint32_t M = h->M, G = h->G, D = h->D, C = h->C;
    if (G == 0 || M == 0)
        return GMM_OK;
    const char* env = getenv("GMM_TC_SWIZZLE");
    tm.swz          = (env && atoi(env) == 64) ? 64 : 128;
    tm.pooled       = (C == 1);
    const uint32_t kel = tm.pooled ? D : 2 * D;
    tm.kSteps          = (kel + 3 + 15) / 16;  // 3 spare K columns carry the per-Gaussian constant
    tm.nAtoms          = (tm.kSteps * 64 + tm.swz - 1) / tm.swz;
    tm.aTileBytes      = tm.nAtoms * kTileM * tm.swz;
    tm.tileBytes       = tm.nAtoms * kTileN * tm.swz + kMetaBytes;

    // ---- column assignment ----
    struct Col {
        uint32_t g;     // slot whose operand row fills the column, or 0xFFFFFFFF for a zero filler column
        uint16_t slot;  // density index within the mixture (0xFFFF: pad / filler)
    };
    std::vector<Col>      cols;
    std::vector<TileMeta> metas;
    std::vector<uint32_t> mixPerHalf;
    auto halfOf = [&]() { return (uint32_t)(cols.size() / 128); };
    auto ensureHalf = [&](uint32_t hf) {
        while (metas.size() * 2 <= hf) {
            TileMeta z = {0, 0, 0, 0, {0, 0, 0, 0}};
            metas.push_back(z);
        }
        while (mixPerHalf.size() <= hf)
            mixPerHalf.push_back(0);
    };
    for (uint32_t m = 0; m < M; ++m) {
        std::vector<uint32_t> live;
        for (uint32_t g = offsets[m]; g < offsets[m + 1]; ++g)
            if (isfinite(w2[g]) && isfinite(logNorm[covIdx[g]]))
                live.push_back(g);
        uint32_t padded = std::max<uint32_t>(8, (uint32_t)((live.size() + 7) / 8 * 8));
        if (padded > 128 || offsets[m + 1] - offsets[m] > 0xFFFF)
            return GMM_OK;  // does not fit the tensor kernel: tm.valid stays false
        const uint32_t inHalf = (uint32_t)(cols.size() % 128);
        if (inHalf + padded > 128) {
            const Col pad = {0xFFFFFFFFu, 0xFFFF};
            cols.insert(cols.end(), 128 - inHalf, pad);
        }
        const uint32_t hf = halfOf();
        ensureHalf(hf);
        TileMeta& mt = metas[hf / 2];
        if (mixPerHalf[hf] == 0)
            (hf & 1 ? mt.mixBase1 : mt.mixBase0) = m;
        mixPerHalf[hf]++;
        const uint32_t seg0 = (uint32_t)((cols.size() % 256) / 8);
        for (uint32_t i = 0; i < padded; ++i) {
            Col c = {0xFFFFFFFFu, 0xFFFF};
            if (i < live.size()) {
                c.g    = live[i];
                c.slot = (uint16_t)(live[i] - offsets[m]);
            }
            else if (!live.empty()) {
                c.g = live[0];  // pad: repeats the first live column, so it never changes a min / strict arg-min
            }
            cols.push_back(c);
        }
        const uint32_t nseg = padded / 8;
        for (uint32_t s = 0; s < nseg; ++s) {
            const uint32_t gs = seg0 + s;
            const uint32_t lv = live.size() > s * 8 ? std::min<uint32_t>(8, (uint32_t)live.size() - s * 8) : 1;
            mt.live[gs >> 3] |= (lv - 1) << ((gs & 7) * 4);
        }
        if (live.empty())
            mt.segEmpty |= 1u << (seg0 + nseg - 1);
        mt.segEnd |= 1u << (seg0 + nseg - 1);
    }
    {
        const Col pad = {0xFFFFFFFFu, 0xFFFF};
        while (cols.size() % kTileN)
            cols.push_back(pad);
    }
    tm.nTiles = (uint32_t)(cols.size() / kTileN);
    ensureHalf(tm.nTiles * 2 - 1);
    tm.maxMixPerHalf = 1;
    for (uint32_t v : mixPerHalf)
        tm.maxMixPerHalf = std::max(tm.maxMixPerHalf, v);

    // ---- shared-memory configuration (sized for the arg-min variant) ----
    const uint32_t pitch = tm.maxMixPerHalf + 1;
    tm.buffersB = 0;
    for (uint32_t nb = 2; nb >= 1 && !tm.buffersB; --nb)
        for (uint32_t sa = kMaxStagesA; sa >= 2; --sa)
            if (smemNeed(tm.tileBytes, tm.aTileBytes, nb, sa, pitch, true) <= kSmemLimit) {
                tm.buffersB = nb;
                tm.stagesA  = sa;
                break;
            }
    if (!tm.buffersB)
        return GMM_OK;  // operand tiles too large for shared memory (very high dimension): direct kernel

    // ---- centre / scale ----
    std::vector<double> ctr(D, 0.0);
    size_t              nLive = 0;
    for (const Col& c : cols)
        if (c.g != 0xFFFFFFFFu) {
            for (uint32_t d = 0; d < D; ++d)
                ctr[d] += mean[(size_t)c.g * D + d];
            ++nLive;
        }
    std::vector<float> ctrF(D), scaleF(D);
    for (uint32_t d = 0; d < D; ++d)
        ctrF[d] = nLive ? (float)(ctr[d] / (double)nLive) : 0.f;
    if (tm.pooled) {
        for (uint32_t d = 0; d < D; ++d)
            scaleF[d] = isr[d];
    }
    else {
        for (uint32_t d = 0; d < D; ++d) {
            double s = 0.0;
            for (uint32_t c = 0; c < C; ++c)
                s += (double)isr[(size_t)c * D + d] * isr[(size_t)c * D + d];
            scaleF[d] = (float)sqrt(s / C);
        }
    }

    // ---- tile images ----
    std::vector<uint8_t> img((size_t)tm.nTiles * tm.tileBytes, 0);
    const uint32_t       epa = tm.swz / 2;
    std::vector<float>   row(tm.kSteps * 16);
    double               kappa = 0.0;  // largest squared Mahalanobis norm of a centred mean
    bool                 cRange = true;
    for (uint32_t t = 0; t < tm.nTiles; ++t) {
        uint8_t*  base  = img.data() + (size_t)t * tm.tileBytes;
        uint16_t* slOut = reinterpret_cast<uint16_t*>(base + (size_t)tm.nAtoms * kTileN * tm.swz);
        TileMeta* mOut  = reinterpret_cast<TileMeta*>(reinterpret_cast<uint8_t*>(slOut) + kTileN * 2);
        *mOut           = metas[t];
        for (uint32_t r = 0; r < (uint32_t)kTileN; ++r) {
            const Col& col = cols[(size_t)t * kTileN + r];
            slOut[r]       = col.slot;
            if (col.g == 0xFFFFFFFFu)
                continue;  // filler / empty mixture: zero operand row
            const float* mu = mean + (size_t)col.g * D;
            const float* is = isr + (size_t)covIdx[col.g] * D;
            std::fill(row.begin(), row.end(), 0.f);
            double cc = (double)logNorm[covIdx[col.g]] + (double)w2[col.g], kap = 0.0;
            for (uint32_t d = 0; d < D; ++d) {
                const double mc = (double)mu[d] - (double)ctrF[d];
                if (tm.pooled) {
                    const double b = mc * (double)is[d];
                    row[d]         = (float)(-2.0 * b);
                    kap += b * b;
                }
                else {
                    const double iv = (double)is[d] * (double)is[d], s = scaleF[d];
                    row[d]          = (float)(iv / (s * s));
                    row[D + d]      = (float)(-2.0 * mc * iv / s);
                    kap += mc * mc * iv;
                }
            }
            kappa = std::max(kappa, kap);
            // constant c = |b|^2 + L + w2 as three fp16 pieces in the spare K columns (the A side holds 1/rowScale)
            const float cf = (float)(cc + kap);
            if (!(fabsf(cf) < 60000.f))
                cRange = false;
            {
                uint16_t h0, h1, h2, dummy;
                host_split(cf, h0, h1);
                __half hh0, hh1;
                memcpy(&hh0, &h0, 2);
                memcpy(&hh1, &h1, 2);
                const float rest = (cf - __half2float(hh0)) - __half2float(hh1);
                host_split(rest, h2, dummy);
                __half hh2;
                memcpy(&hh2, &h2, 2);
                row[kel]     = __half2float(hh0);
                row[kel + 1] = __half2float(hh1);
                row[kel + 2] = __half2float(hh2);
            }
            for (uint32_t k = 0; k < tm.kSteps * 16; ++k) {
                uint16_t hi, lo;
                host_split(row[k], hi, lo);
                const uint32_t kl = tm.kSteps * 16 + k;
                uint8_t* pHi = base + (size_t)(k / epa) * kTileN * tm.swz + swz_offset(r, (k % epa) / 8, tm.swz) + (k % 8) * 2;
                uint8_t* pLo = base + (size_t)(kl / epa) * kTileN * tm.swz + swz_offset(r, (kl % epa) / 8, tm.swz) + (kl % 8) * 2;
                memcpy(pHi, &hi, 2);
                memcpy(pLo, &lo, 2);
            }
        }
    }
    // The expanded form |a|^2 + |b|^2 - 2 a.b cancels in f32: its absolute error grows like 2^-22 * |b|^2 while a
    // frame next to such a mean has a score of order D/2.  Models whose centred, whitened means are further out
    // than 8 sigma (rms per dimension) keep the 1e-4 relative bar only in the direct kernel.
    if (!force && kappa > 64.0 * D)
        return GMM_OK;
    if (!cRange) {
        if (force) {
            h->setError("tensor kernel: a per-Gaussian constant exceeds the fp16 range");
            return GMM_ERR_UNSUPPORTED;
        }
        return GMM_OK;
    }
    int rc;
    if ((rc = h->upload((void**)&tm.bTiles, img.data(), img.size())))
        return rc;
    if ((rc = h->upload((void**)&tm.center, ctrF.data(), D * 4)))
        return rc;
    if ((rc = h->upload((void**)&tm.scale, scaleF.data(), D * 4)))
        return rc;
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    tm.valid = true;
    return GMM_OK;
}

int gmm_tc_score(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld) {
    const TensorModel& tm = h->tensor;
    if (T == 0)
        return GMM_OK;
    const uint32_t FT = (T + kTileM - 1) / kTileM;
    void *         aT = nullptr, *rT = nullptr;
    int            rc;
    if ((rc = h->reserve(SCR_A, (size_t)FT * tm.aTileBytes, &aT)))
        return rc;
    if ((rc = h->reserve(SCR_ROW, (size_t)FT * kTileM * 8, &rT)))
        return rc;
    {
        int blocks = (int)std::min<uint32_t>((FT * kTileM + 7) / 8, (uint32_t)h->smCount * 8);
        if (tm.pooled)
            tc_prep_kernel<true><<<blocks, 256, 0, h->stream>>>(d_x, T, h->D, FT, tm.center, tm.scale, tm.kSteps,
                                                                tm.nAtoms, tm.swz, (uint8_t*)aT, (float2*)rT);
        else
            tc_prep_kernel<false><<<blocks, 256, 0, h->stream>>>(d_x, T, h->D, FT, tm.center, tm.scale, tm.kSteps,
                                                                 tm.nAtoms, tm.swz, (uint8_t*)aT, (float2*)rT);
        h->launches++;
        GMM_CUDA_TRY(h, cudaGetLastError());
    }
    KParams p;
    p.bTiles     = tm.bTiles;
    p.aTiles     = (const uint8_t*)aT;
    p.rowInfo    = (const float2*)rT;
    p.scores     = d_scores;
    p.best       = d_best;
    p.ld         = ld;
    p.T          = T;
    p.nTiles     = tm.nTiles;
    p.frameTiles = FT;
    p.kSteps     = tm.kSteps;
    p.atoms      = tm.nAtoms;
    p.aTileBytes = tm.aTileBytes;
    p.bTileBytes = tm.tileBytes;
    p.buffersB   = tm.buffersB;
    p.stagePitch = tm.maxMixPerHalf + 1;
    const bool best = d_best != nullptr;
    // A ring depth: whatever shared memory allows for this variant
    p.stagesA = tm.stagesA;
    while (p.stagesA < (uint32_t)kMaxStagesA &&
           smemNeed(tm.tileBytes, tm.aTileBytes, tm.buffersB, p.stagesA + 1, p.stagePitch, best) <= kSmemLimit)
        p.stagesA++;
    const size_t smemBytes = smemNeed(tm.tileBytes, tm.aTileBytes, tm.buffersB, p.stagesA, p.stagePitch, best);
    // split the frame range so that the items fill the SMs evenly: cost in A-tile MMA times
    const double bCost = tm.buffersB == 2 ? 0.5 : 2.0;
    uint32_t     bestS = 1;
    double       bestCost = 1e300;
    for (uint32_t S = 1; S <= std::min<uint32_t>(FT, 64); ++S) {
        const uint32_t per = (FT + S - 1) / S;
        if ((uint64_t)per * (S - 1) >= FT)
            continue;  // an empty chunk
        const uint64_t items  = (uint64_t)tm.nTiles * S;
        const uint64_t rounds = (items + h->smCount - 1) / h->smCount;
        const double   cost   = (double)rounds * (per + bCost);
        if (cost < bestCost - 1e-9) {
            bestCost = cost;
            bestS    = S;
        }
    }
    p.split        = bestS;
    const int grid = (int)std::min<uint64_t>((uint64_t)h->smCount, (uint64_t)tm.nTiles * bestS);
    cudaError_t e;
    const int   mode = h->mode == GMM_MODE_SUM ? 1 : 0;
#define GMM_TC_LAUNCH(SWZ)                                                                                   \
    (mode == 0 ? (best ? launch_tc<SWZ, 0, true>(p, grid, smemBytes, h->stream)                               \
                       : launch_tc<SWZ, 0, false>(p, grid, smemBytes, h->stream))                             \
               : (best ? launch_tc<SWZ, 1, true>(p, grid, smemBytes, h->stream)                               \
                       : launch_tc<SWZ, 1, false>(p, grid, smemBytes, h->stream)))
    if (tm.swz == 128)
        e = GMM_TC_LAUNCH(128);
    else
        e = GMM_TC_LAUNCH(64);
#undef GMM_TC_LAUNCH
    h->launches++;
    GMM_CUDA_TRY(h, e);
    return GMM_OK;
}
