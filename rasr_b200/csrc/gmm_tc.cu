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
// repeats the mixture's first live column); a mixture never straddles a column part (the 128 or 64 columns one
// epilogue warp reduces).  TileMeta: bit s of segEnd says "8-column segment s closes a mixture", plus live-column
// counts, the first mixture per part and the uniform-shape flags the epilogue specialises on.
//
// Kernel: persistent, one CTA per SM.  Warp 0: bulk-copy producer (one lane) and TMEM allocation; warp 1: MMA
// issuer (warp converged, one elected lane, M=128 N=256 K=16 per instruction, descriptors in uniform registers);
// then 8 or 16 epilogue warps (TMEM lane quarter = warp % 4, column part = (warp - 2) / 4).  Work item = (B tile,
// range of frame tiles); B stays in shared memory for the whole item while A tiles stream through a ring; two
// 256-column accumulators in TMEM double-buffer MMA against the epilogue.  Synchronisation: tile-indexed mbarrier
// rings ready[] / done[] (see tc_score_body).  DESIGN.md 4.1 lists the measurements behind these choices.
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <type_traits>

#include "gmm_internal.cuh"
#include "gmm_umma.cuh"

namespace {

// The 6-instruction operand split: the correction operands are e4m3 (4 significant bits).  A lo part is ~2^-12 of its
// value, far below e4m3's range next to the hi parts it multiplies, so lo parts are stored times 2^kLoShift: the e4m3
// instructions leave 2^kLoShift * (x_lo.mu_hi + x_hi.mu_lo) in the accumulator and the FIRST fp16 instruction, issued
// after them, scales the accumulator input back (tcgen05.mma scale-input-d: D = A.B + D * 2^-kLoShift, exact).
constexpr int   kLoShift = 8;
constexpr float kLoScale = 256.f;

struct KParams {
    const uint8_t* bTiles;
    const uint8_t* aTiles;
    const float2*  rowInfo;  // per frame row: {row term |a|^2, power-of-two row scale}
    float*         scores;
    uint32_t*      best;
    size_t         ld;
    uint32_t       T;
    uint32_t       nTiles, frameTiles, split;  // items = nTiles * split
    uint32_t       swapTile;                   // B tile that trades places with the last one (nTiles - 1: none), see launch
    uint32_t       kSteps;                     // k-steps (16 elements) per precision part
    uint32_t       fp8;                        // part 2 holds e4m3 corrections (TensorModel::fp8)
    uint32_t       atoms;                      // swizzle atoms per operand row
    uint32_t       aTileBytes, bTileBytes;     // bTileBytes includes c / slot / meta
    uint32_t       stagesA, buffersB;
    uint32_t       stagePitch;                 // floats per staging row (odd, >= maxMixPerPart + 5)
    const uint32_t* itemTab;                   // subset scoring: 8 words per item {B tile, f0, f1, rowBase, rowEnd}
    uint32_t       nItems;
    unsigned long long* dbg;                   // GMM_TC_DEBUG: per-CTA cycle counters of the MMA thread
    uint32_t       directStores;               // 0: always stage and store coalesced (GMM_TC_STORES=staged)
    uint32_t       sumFull;                    // full-sum scorer: parts without dead columns take the unpredicated pass
};

// exp2 on the FMA and ALU pipes for arguments <= 0: with -DGMM_SUM_POLY_EVERY=n every n-th term of the full-sum
// epilogue takes this way instead of the special-function unit.  Measured on the final (spill-free) kernel, 12001x64 /
// 4501x16: n = 8: 2.31 -> 2.27 ms / 0.767 -> 0.747 ms; n = 4: 2.42 / 0.758 ms -- two percent at best, so the default
// stays 0 (every term through ex2.approx, one error model).  Round to the nearest integer n with the 1.5 * 2^23 trick, degree-5 polynomial
// of 2^f on f = t - n in [-0.5, 0.5] (Taylor in f ln 2: relative error <= 2.4e-6), n added to the exponent field.
#ifndef GMM_SUM_POLY_EVERY
#define GMM_SUM_POLY_EVERY 0
#endif
constexpr int kSumPolyEvery = GMM_SUM_POLY_EVERY;
__device__ __forceinline__ float exp2_fma(float t) {
    t             = fmaxf(t, -126.f);
    const float r = __fadd_rn(t, 12582912.f);
    const float f = __fsub_rn(t, __fsub_rn(r, 12582912.f));
    float       p = fmaf(f, 1.3333558146e-3f, 9.6181291076e-3f);
    p             = fmaf(p, f, 5.5504108665e-2f);
    p             = fmaf(p, f, 2.4022650696e-1f);
    p             = fmaf(p, f, 6.9314718056e-1f);
    p             = fmaf(p, f, 1.0f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(r) << 23));
}

struct EpiState {
    float    run, sum;
    uint32_t idx;
    uint32_t sp;  // shared-memory address of this row's next staged score (the index plane lies offB behind)
};

// One 32-column chunk (four 8-column segments) of a row's accumulator.  The accumulator already holds
// u / rowScale with u = c - 2 a.b (the constant rides in spare K columns of the contraction), so the maximum
// approximation is a bare min tree over TMEM values; a closing segment stages the mixture's score.
// SP > 0: every mixture of this column part spans exactly SP segments and none is empty (a property of the packed
// tile, TileMeta::uni), so the closing positions are compile-time constants; SP = 0: driven by the segEnd bits.
// res != nullptr (SP > 0, maximum approximation without arg-min): the mixture scores stay in registers, res[seg / SP],
// and the caller stores the row piece straight from them (no staging).
template<int MODE, bool BEST, int SP>
__device__ __forceinline__ void epi_chunk(const uint32_t* v, int ch, int colBase, const uint16_t* __restrict__ slots,
                                          uint32_t segEnd, uint32_t segEmpty, uint64_t liveBits, float rt, float rs,
                                          uint32_t offB, EpiState& st, float* res = nullptr) {
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
            // three-input min tree; a closing segment stores 0.5 * (min * rowScale + rowTerm)
            // = fma(min, 0.5 rowScale, 0.5 rowTerm) (rowScale is a power of two, so both roundings agree)
            // four 3-input minima per segment: (u0 u1 u2) (u3 u4 u5) (u6 u7 run) and the three results
            const float ma = fminf(fminf(u[0], u[1]), u[2]), mb = fminf(fminf(u[3], u[4]), u[5]);
            if (SP > 0) {
                const bool  first = (seg % SP) == 0, last = (seg % SP) == SP - 1;  // compile-time after unrolling
                const float mc = first ? fminf(u[6], u[7]) : fminf(fminf(u[6], u[7]), st.run);
                const float m  = fminf(fminf(ma, mb), mc);
                if (last) {
                    const float sc = fminf(fmaf(m, 0.5f * rs, 0.5f * rt), 0.5f * (FLT_MAX + rt));
                    if (res)
                        res[seg / SP] = sc;
                    else {
                        sts_f32(st.sp, sc);
                        st.sp += 4;
                    }
                }
                st.run = m;
                continue;
            }
            const float m = fminf(fminf(ma, mb), fminf(fminf(u[6], u[7]), st.run));
            const bool  close = (segEnd >> seg) & 1u;
            float       sc    = fminf(fmaf(m, 0.5f * rs, 0.5f * rt), 0.5f * (FLT_MAX + rt));
            sc                = ((segEmpty >> seg) & 1u) ? 0.5f * (FLT_MAX + rt) : sc;
            if (close)
                sts_f32(st.sp, sc);
            st.sp += close ? 4u : 0u;
            st.run = close ? INFINITY : m;
            continue;
        }
        else if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (u[j] < st.run) {  // pads repeat the first live column: never strictly smaller
                    st.run = u[j];
                    st.idx = (uint32_t)(colBase + seg * 8 + j);
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
                    st.idx = (uint32_t)(colBase + seg * 8 + j8);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (j < live)
                        st.sum += ex2((st.run - u[j]) * hl2e);
            }
        }
        if (SP > 0 ? (seg % SP) == SP - 1 : (bool)((segEnd >> seg) & 1u)) {  // uniform across the warp
            const bool  empty = SP > 0 ? false : (bool)((segEmpty >> seg) & 1u);
            const float best  = empty ? FLT_MAX : fminf(st.run * rs, FLT_MAX);
            float       sc    = 0.5f * (best + rt);
            if (MODE == 1)
                sc = empty ? INFINITY : sc - kLn2 * __log2f(st.sum);
            sts_f32(st.sp, sc);
            if (BEST)
                sts_u32(st.sp + offB, (empty || st.idx == 0xFFFFFFFFu) ? 0xFFFFFFFFu : (uint32_t)slots[st.idx]);
            st.sp += 4;
            st.run = INFINITY;
            st.sum = 0.f;
            st.idx = 0xFFFFFFFFu;
        }
    }
}

// Full-sum scorer on a column part whose mixtures all span SP segments (TileMeta::uni): the mixture's values are all
// in this thread's registers, so the log-sum-exp is two passes -- the minimum first (a min tree, as cheap as the
// maximum approximation), then one exp2 per live density against that fixed reference -- instead of the online form
// with its running-minimum updates and rescaling.  A segment whose densities lie, for every frame of the warp, more
// than 40 above the mixture's best is skipped: exp(-40 / 2) < 2^-28 cannot change an f32 sum that is >= 1.
// Measured and dropped: moving a share of the exponentials from the special-function unit to a polynomial on the FMA
// pipe (12001x64 model, 10000 frames: 2.91 ms with none, 3.83 / 4.88 / 5.93 ms with 2 / 4 / 6 of 8) -- the epilogue is
// bound by instruction issue, not by the 16 exp2 per clock of the SFU.
// Also measured and dropped (round 2): a branch for full segments with eight unpredicated fma -> ex2 pairs and an add
// tree instead of the predicated serial chain below -- 3.53 -> 4.16 ms (eight more live registers at the 96 the
// 16-warp kernel has).
// The same tree with EIGHT epilogue warps (168 registers, the part's 128 columns as two 64-column halves): 4.54 ms -- two
// warps per scheduler do not cover the special-function latency, sixteen do.
// And in place (arguments, then exponentials, then adds, each value overwriting its own register: no extra registers):
// 3.71 -> 4.08 ms.  The serial chain interleaves best with the other three warps of the scheduler.  What the data say
// (12001x64 synthetic model): a frame needs 41 % of its segments at the threshold of 40, a WARP of 32 independent frames
// 99.99 % -- the vote skips nothing, the kernel pays all 64 exponentials per frame and mixture (>= 2048 cycles per tile).
// get(i) = accumulator value i of the part (compile-time i), N = columns of the part held in registers.
// FULL: every column of the part is a live density (liveBits all 7) -- no predicates, no vote: the vote skips nothing on
// independent frames (above), and without it and the `j < live` tests the loop is multiply-add, exp2, add per term
// (12001x64 model, 10000 frames: 3.98 -> 3.32 ms in the same build; GMM_TC_SUM_FULL=0 switches it off).
// Measured around this form at the end of round 2 (each as its own build, profiles/sum_epilogue_experiments_r02.txt):
//   * ex2 replaced by the identity (timing only): 2.67 ms -- the special-function unit accounts for a fifth of the tile,
//     not for the 2048 of its ~3500 cycles the 16-results-per-clock arithmetic suggested;
//   * fma.rn.f32x2 / add.rn.f32x2 (FFMA2 / FADD2, half the FMA-pipe instructions): 3.56 ms, and 2.96 ms without ex2 --
//     slower both times: neither the FMA pipe nor instruction issue is the bound;
//   * 32 accumulator values live instead of 64 (mixtures of <= 32 columns chunk by chunk; 64-column mixtures with the
//     first chunk read from TMEM twice): 3.85 ms;
//   * the four epilogue warps of a scheduler started 300 .. 900 cycles apart: no change;
//   * an event log of one CTA: the tensor core finishes a tile 530 cycles after its commit and runs two tiles ahead of
//     the epilogue the whole time -- the epilogue warps never wait for it.
// What remains is the per-warp dependency chain of a tile (TMEM load, min tree, 64 dependent adds behind 64 exp2
// results, log2, staged store), which four warps per scheduler do not overlap: the next step is more, narrower
// warps' worth of independent chains per scheduler, i.e. two mixtures' sums interleaved in one thread.
template<bool BEST, int SP, int NSEG, bool FULL, typename Get>
__device__ __forceinline__ void epi_sum_uniform(Get get, int seg0, int colBase, const uint16_t* __restrict__ slots,
                                                uint64_t liveBits, float rt, float rs, uint32_t offB, EpiState& st,
                                                float* res = nullptr) {
    constexpr float kLn2 = 0.69314718055994530942f;
    const float     hl2e = 0.72134752044448170368f * rs;  // 0.5 * log2(e) * rowScale
#pragma unroll
    for (int m0 = 0; m0 < NSEG; m0 += SP) {  // one mixture: segments [m0, m0 + SP) of the registers
        float segMin[SP];
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            const int b = (m0 + s) * 8;
            segMin[s]   = fminf(fminf(fminf(get(b), get(b + 1)), fminf(get(b + 2), get(b + 3))),
                                fminf(fminf(get(b + 4), get(b + 5)), fminf(get(b + 6), get(b + 7))));
        }
        float best = segMin[0];
#pragma unroll
        for (int s = 1; s < SP; ++s)
            best = fminf(best, segMin[s]);
        const float bestL = best * hl2e;
        float       sum   = 0.f;
        if (FULL) {
#pragma unroll
            for (int i = 0; i < SP * 8; ++i) {
                const float t = fmaf(get(m0 * 8 + i), -hl2e, bestL);
                sum += (kSumPolyEvery > 0 && i % (kSumPolyEvery > 0 ? kSumPolyEvery : 1) == kSumPolyEvery / 2) ? exp2_fma(t) : ex2(t);
            }
        }
        else {
#pragma unroll
            for (int s = 0; s < SP; ++s) {
                const int live = (int)((liveBits >> ((seg0 + m0 + s) * 4)) & 15ull) + 1;  // uniform
                if (__any_sync(0xffffffffu, !((segMin[s] - best) * rs >= 40.f))) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (j < live)  // (best - u) * hl2e as one fused multiply-add: the epilogue is issue-bound
                            sum += ex2(fmaf(get((m0 + s) * 8 + j), -hl2e, bestL));
                }
            }
        }
        const float b2 = fminf(best * rs, FLT_MAX);
        const float sc = 0.5f * (b2 + rt) - kLn2 * __log2f(sum);
        if (res)  // the caller writes the row piece from registers (as the maximum approximation's direct stores)
            res[m0 / SP] = sc;
        else
            sts_f32(st.sp, sc);
        if (BEST) {
            uint32_t idx = 0;
#pragma unroll
            for (int i = SP * 8 - 1; i >= 0; --i)  // first density that reaches the minimum
                idx = get(m0 * 8 + i) == best ? (uint32_t)i : idx;
            sts_u32(st.sp + offB, (uint32_t)slots[colBase + (seg0 + m0) * 8 + idx]);
        }
        st.sp += 4;
    }
}

// MODE 0 = max, 1 = sum.  BEST: also write the index of the winning density.
// NE = epilogue warps (8 or 16): warp e handles TMEM lane quarter (warp & 3) and column part e / 4 of
// 1024 / NE columns (128 or 64).  16 warps give every scheduler four epilogue warps to hide the latencies of
// the min tree, and drain an accumulator in half the time; they need mixtures of at most 64 packed columns.
// Work item -> B tile, frame tiles [f0, f1), and where the rows of frame tile f0 land in the score block.
// Full scoring: items enumerate (B tile, chunk of the frame range), rows are the frames themselves.  Subset scoring
// (gmm_tc_score_subset): an item table pairs a segment's own B tile with the segment's frame tiles, whose rows are
// the segment's frames [rowBase, rowEnd) although its A tiles sit in a tile-aligned (padded) row space.
struct ItemRange {
    uint32_t nt, f0, f1, rowBase, rowEnd;
};
__device__ __forceinline__ ItemRange item_range(const KParams& p, uint32_t item, uint32_t tilesPerItem) {
    ItemRange r;
    if (p.itemTab) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(p.itemTab) + 2 * (size_t)item);
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(p.itemTab) + 2 * (size_t)item + 1);
        r.nt = a.x, r.f0 = a.y, r.f1 = a.z, r.rowBase = a.w, r.rowEnd = b.x;
    }
    else {
        const uint32_t chunk = item % p.split;
        r.nt      = item / p.split;
        r.nt      = r.nt == p.nTiles - 1 ? p.swapTile : r.nt == p.swapTile ? p.nTiles - 1 : r.nt;
        r.f0      = chunk * tilesPerItem;
        r.f1      = min(p.frameTiles, r.f0 + tilesPerItem);
        r.rowBase = r.f0 * kTileM;
        r.rowEnd  = p.T;
    }
    return r;
}

// NG = 2 (NE = 8 only): TWO epilogue groups of eight warps; group g drains the accumulator of the tiles with t % 2 == g.
// With one group the epilogue's phases alternate -- TMEM loads and the min tree (ALU pipe), then the row pieces (load /
// store unit, memory latency) -- and the MMA warp waits for the accumulator a third of the time (4501x16 model: 337 of
// 1189 cycles per tile); with two, one group's stores run under the other group's reduction.
template<int SWZ, int MODE, bool BEST, int NE, int NG = 1>
__device__ __forceinline__ void tc_score_body(const KParams& p) {
    static_assert(NG == 1 || (NG == 2 && NE == 8 && kRing % 2 == 0), "two groups: eight warps each, ring parity = tile parity");
    // First epilogue warp.  Sixteen epilogue warps in one group: the CTA is launched with TWENTY warps -- registers are
    // handed out per four warps, so eighteen warps pay for twenty anyway (96 registers each) -- and the first warp group
    // (producer, MMA issuer, two idle warps) gives registers back: setmaxnreg 32 there, 112 for the epilogue warps, at
    // which the full-sum epilogue (64 accumulator values live per thread) no longer spills (ptxas: 228 bytes at 96).
    // (Full-sum scorer without arg-min only: the other sixteen-warp variants need more than 112 registers somewhere, and
    // a setmaxnreg region cannot spill -- ptxas refuses.)
    constexpr int EW0 = (NE == 16 && NG == 1 && !BEST) ? 4 : kEpiWarp0;
    extern __shared__ uint8_t smemRaw[];
    // swizzled operand tiles need 1024-byte alignment of the shared-memory address itself
    uint8_t*  smem = smemRaw + ((1024u - (smem_u32(smemRaw) & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int PW   = 1024 / NE;  // columns per epilogue warp
    constexpr int SEGS = PW / 8;     // 8-column segments per epilogue warp

    // carve shared memory: [B buffers][A stages][staging][barriers][MMA descriptor offsets]
    const uint32_t bBytes    = (p.bTileBytes + 1023u) & ~1023u;
    uint8_t*       sB        = smem;
    uint8_t*       sA        = sB + (size_t)p.buffersB * bBytes;
    float*         sStage    = reinterpret_cast<float*>(sA + (size_t)p.stagesA * p.aTileBytes);
    const uint32_t stageWarp = 32 * p.stagePitch;  // floats per warp
    uint32_t*      sStageB   = reinterpret_cast<uint32_t*>(sStage + NE * NG * stageWarp);
    uint64_t*      bars      = reinterpret_cast<uint64_t*>(
            (reinterpret_cast<uintptr_t>(sStageB + (BEST ? NE * NG * stageWarp : 0)) + 15) & ~(uintptr_t)15);
    // Barriers.  Tiles are numbered t = 0, 1, ... per CTA across its items; rings of kRing follow t % kRing.
    //   ready[t % 4]: tile t may be issued = its A operand has landed (producer arrive + bytes) AND the
    //                 accumulator t & 1 was drained by every epilogue warp (after tile t - 2)
    //   done[t % 4] : the MMAs of tile t have retired (one tcgen05.commit): the epilogue may read the
    //                 accumulator, the producer may overwrite the A stage
    uint64_t* ready    = bars;                // [kRing]
    uint64_t* done     = bars + kRing;        // [kRing]
    uint64_t* fullB    = bars + 2 * kRing;    // [2]
    uint64_t* emptyB   = fullB + 2;           // [2]
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(emptyB + 2);

    if (threadIdx.x == 32) {
        for (int i = 0; i < kRing; ++i) {
            mbar_init(smem_u32(&ready[i]), 1 + NE);
            mbar_init(smem_u32(&done[i]), 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(smem_u32(&fullB[i]), 1);
            mbar_init(smem_u32(&emptyB[i]), 1 + NE * NG);  // MMA commit + the epilogue warps
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

    const uint32_t items        = p.itemTab ? p.nItems : p.nTiles * p.split;
    const uint32_t tilesPerItem = (p.frameTiles + p.split - 1) / p.split;

    if (warp < EW0) {
    if (EW0 == 4)
        asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
    if (warp == 0) {
        // ===== producer: one lane issues the bulk copies =====
        if (lane == 0) {
            uint32_t itemCount = 0, t = 0, stage = 0;
            for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
                const ItemRange ir = item_range(p, item, tilesPerItem);
                const uint32_t  nt = ir.nt, f0 = ir.f0, f1 = ir.f1;
                if (f0 >= f1)
                    continue;
                const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
                mbar_wait(smem_u32(&emptyB[b]), (useB & 1) ^ 1);
                mbar_expect_tx(smem_u32(&fullB[b]), p.bTileBytes);
                bulk_g2s(smem_u32(sB + (size_t)b * bBytes), p.bTiles + (size_t)nt * p.bTileBytes, p.bTileBytes,
                         smem_u32(&fullB[b]));
                for (uint32_t f = f0; f < f1; ++f, ++t) {
                    if (t >= p.stagesA) {  // the stage's previous tenant has been consumed by the tensor core
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
        // ===== MMA issuer: the warp stays converged (every operand below is warp-uniform, so descriptors live in
        // uniform registers); one elected lane issues tcgen05.mma / tcgen05.commit for the whole CTA =====
        // instruction descriptor: D=f32, A=B=f16, K-major both, N=256, M=128
        constexpr uint32_t idesc    = (1u << 4) | ((uint32_t)(kTileN >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
        constexpr uint32_t kPerAtom = SWZ / 32;  // 16-element k-steps per swizzle atom
        // high word of the shared-memory descriptor: SBO = 8 rows * SWZ bytes (>> 4), version 1, swizzle mode
        constexpr uint32_t descHi   = (uint32_t)((8 * SWZ) >> 4) | (1u << 14) | ((SWZ == 128 ? 2u : 4u) << 29);
        const bool         leader   = elect_one();
        const uint32_t     kS       = p.kSteps;
        uint32_t  itemCount = 0, t = 0, stage = 0;
        bool      have = false;  // ready[t % kRing] was already seen complete (probed while tile t - 1 was being issued)
        long long tStart = clock64(), wR = 0, wI = 0, t0 = 0, wB = 0;
        unsigned long long gt0 = 0;
        if (p.dbg) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt0));
        for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
            const ItemRange ir = item_range(p, item, tilesPerItem);
            const uint32_t  f0 = ir.f0, f1 = ir.f1;
            if (f0 >= f1)
                continue;
            const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
            if (p.dbg) t0 = clock64();
            mbar_wait(smem_u32(&fullB[b]), useB & 1);
            if (p.dbg) wB += clock64() - t0;
            const uint32_t bLo = (smem_u32(sB + (size_t)b * bBytes) & 0x3FFFF) >> 4;
            for (uint32_t f = f0; f < f1; ++f, ++t) {
                if (p.dbg) t0 = clock64();
                if (!have)
                    mbar_wait(smem_u32(&ready[t % kRing]), (t / kRing) & 1);
                tc_fence_after();
                if (p.dbg) { const long long t1 = clock64(); wR += t1 - t0; t0 = t1; }
                const uint32_t aLo   = (smem_u32(sA + (size_t)stage * p.aTileBytes) & 0x3FFFF) >> 4;
                const uint32_t dTmem = tmemBase + (t & 1) * kTileN;
                if (leader && p.fp8) {
                    // first ONE e4m3 instruction per k-step over part 2 (2^8 (x_lo.mu_hi + x_hi.mu_lo)), then hi.hi in
                    // fp16, whose first instruction takes the accumulator scaled by 2^-8; the e4m3 formats are code 0 of
                    // kind::f8f6f4, so the instruction descriptor is the fp16 one
                    for (uint32_t ks = 0; ks < kS; ++ks) {
                        const uint32_t kl = kS + ks;
                        const uint32_t al = aLo + (kl / kPerAtom) * (kTileM * SWZ / 16) + (kl % kPerAtom) * 2;
                        const uint32_t bl = bLo + (kl / kPerAtom) * (kTileN * SWZ / 16) + (kl % kPerAtom) * 2;
                        tc_mma_f8(dTmem, ((uint64_t)descHi << 32) | al, ((uint64_t)descHi << 32) | bl, idesc, ks ? 1u : 0u);
                    }
                    for (uint32_t ks = 0; ks < kS; ++ks) {
                        const uint32_t ah = aLo + (ks / kPerAtom) * (kTileM * SWZ / 16) + (ks % kPerAtom) * 2;
                        const uint32_t bh = bLo + (ks / kPerAtom) * (kTileN * SWZ / 16) + (ks % kPerAtom) * 2;
                        if (ks == 0)
                            tc_mma_f16_scaled<kLoShift>(dTmem, ((uint64_t)descHi << 32) | ah, ((uint64_t)descHi << 32) | bh, idesc);
                        else
                            tc_mma_f16(dTmem, ((uint64_t)descHi << 32) | ah, ((uint64_t)descHi << 32) | bh, idesc, 1u);
                    }
                    tc_commit(smem_u32(&done[t % kRing]));
                    if (f + 1 == f1)
                        tc_commit(smem_u32(&emptyB[b]));
                }
                else if (leader) {
                    // per k-step: hi.hi, lo.hi, hi.lo (hi part = k-step ks, lo part = k-step kS + ks of the image)
                    for (uint32_t ks = 0; ks < kS; ++ks) {
                        const uint32_t kl = kS + ks;
                        const uint32_t ah = aLo + (ks / kPerAtom) * (kTileM * SWZ / 16) + (ks % kPerAtom) * 2;
                        const uint32_t al = aLo + (kl / kPerAtom) * (kTileM * SWZ / 16) + (kl % kPerAtom) * 2;
                        const uint32_t bh = bLo + (ks / kPerAtom) * (kTileN * SWZ / 16) + (ks % kPerAtom) * 2;
                        const uint32_t bl = bLo + (kl / kPerAtom) * (kTileN * SWZ / 16) + (kl % kPerAtom) * 2;
                        tc_mma_f16(dTmem, ((uint64_t)descHi << 32) | ah, ((uint64_t)descHi << 32) | bh, idesc, ks ? 1u : 0u);
                        tc_mma_f16(dTmem, ((uint64_t)descHi << 32) | al, ((uint64_t)descHi << 32) | bh, idesc, 1u);
                        tc_mma_f16(dTmem, ((uint64_t)descHi << 32) | ah, ((uint64_t)descHi << 32) | bl, idesc, 1u);
                    }
                    tc_commit(smem_u32(&done[t % kRing]));  // accumulator ready, A stage free
                    if (f + 1 == f1)
                        tc_commit(smem_u32(&emptyB[b]));    // B buffer no longer read by the tensor core
                }
                // probe the next tile's barrier while this tile's MMAs are still queued in the tensor core
                have = mbar_test(smem_u32(&ready[(t + 1) % kRing]), ((t + 1) / kRing) & 1);
                __syncwarp();
                if (++stage == p.stagesA)
                    stage = 0;
                if (p.dbg) wI += clock64() - t0;
            }
        }
        if (p.dbg && leader) {
            unsigned long long* d = p.dbg + blockIdx.x * 128;
            unsigned long long gt1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt1));
            d[0] = clock64() - tStart, d[1] = wR, d[2] = wI, d[3] = wB, d[4] = t, d[5] = itemCount, d[6] = gt0, d[7] = gt1;
        }
    }
    }
    else {
        if (EW0 == 4)
            asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
        // ===== epilogue: TMEM -> registers -> (min / log-sum-exp per mixture) -> staged -> global =====
        const int      grp = (warp - EW0) / NE, e = (warp - EW0) % NE, q = warp & 3, part = e >> 2;
        const uint32_t stWarp = smem_u32(sStage) + (uint32_t)(grp * NE + e) * stageWarp * 4;  // this warp's staging block
        const uint32_t stRow  = stWarp + (uint32_t)lane * p.stagePitch * 4;      // this lane's (frame's) row in it
        const uint32_t offB   = NE * NG * stageWarp * 4;                         // index plane behind the scores
        const ptrdiff_t bestDelta = BEST ? reinterpret_cast<const char*>(p.best) - reinterpret_cast<const char*>(p.scores) : 0;
        uint32_t       itemCount = 0, t = 0;
        long long      eW = 0, eL = 0, eC = 0, eS = 0, ec0 = 0;  // GMM_TC_DEBUG: where this warp's cycles go
        if (lane == 0) {  // both accumulators start out drained (two groups: each speaks for its own)
            if (NG == 1 || grp == 0)
                mbar_arrive(smem_u32(&ready[0]));
            if (NG == 1 || grp == 1)
                mbar_arrive(smem_u32(&ready[1]));
        }
        for (uint32_t item = blockIdx.x; item < items; item += gridDim.x, ++itemCount) {
            const ItemRange ir = item_range(p, item, tilesPerItem);
            const uint32_t  f0 = ir.f0, f1 = ir.f1;
            if (f0 >= f1)
                continue;
            const uint32_t b = itemCount % p.buffersB, useB = itemCount / p.buffersB;
            mbar_wait(smem_u32(&fullB[b]), useB & 1);
            // this group's first tile of the item (tiles alternate between the groups across items as well)
            const uint32_t  fG = NG == 2 ? f0 + ((t ^ (uint32_t)grp) & 1u) : f0;
            const uint32_t  tItem = t;
            const uint8_t*  tile  = sB + (size_t)b * bBytes + (size_t)p.atoms * kTileN * SWZ;
            const uint16_t* slots = reinterpret_cast<const uint16_t*>(tile);
            const uint32_t* meta  = reinterpret_cast<const uint32_t*>(tile + kTileN * 2);  // TileMeta words
            constexpr uint32_t segMask = (1u << SEGS) - 1u;
            const uint32_t  segEnd   = (meta[0] >> (SEGS * part)) & segMask;
            const uint32_t  segEmpty = (meta[1] >> (SEGS * part)) & segMask;
            const uint32_t  mixBase  = meta[2 + part];
            const uint64_t  liveBits = (SEGS == 16) ? ((uint64_t)meta[6 + 2 * part + 1] << 32 | meta[6 + 2 * part])
                                                    : (uint64_t)meta[6 + part];
            // full-sum scorer, 16 warps (8 segments per part): no dead column in the part
            const bool      sumFull  = MODE == 1 && p.sumFull && (uint32_t)liveBits == 0x77777777u;
            const uint32_t  nMix     = __popc(segEnd);
            // staged block [32 rows x nMix] -> global: lane starts at element `lane`, every pass advances 32 elements
            const uint32_t  adv_r = 32 / max(nMix, 1u), adv_j = 32 - adv_r * nMix;
            const uint32_t  lr0 = lane / max(nMix, 1u), lj0 = lane - lr0 * nMix;
            const size_t    gStep = ((size_t)adv_r * p.ld + adv_j) * 4, gStepW = gStep + (p.ld - nMix) * 4;  // bytes
            const uint32_t  sStep = (adv_r * p.stagePitch + adv_j) * 4, sStepW = sStep + (p.stagePitch - nMix) * 4;
            const uint32_t  ssLane = stWarp + (lr0 * p.stagePitch + lj0) * 4;
            char*           gTile  = reinterpret_cast<char*>(p.scores + (size_t)(ir.rowBase + (fG - f0) * kTileM + q * 32 + lr0) * p.ld + mixBase + lj0);
            const size_t    gTileStep = (size_t)NG * kTileM * p.ld * 4;
            const float2*   rip = p.rowInfo + (size_t)fG * kTileM + q * 32 + lane;
            // the tile loop, specialised on the tile's mixture shape when it is uniform (see epi_chunk)
            // DIRECT (no arg-min, uniform parts, row pieces 8- / 16-byte aligned; maximum approximation, and the full-sum scorer
            // with its 16 warps): a lane
            // keeps its frame's R = SEGS / SP mixture scores in registers and writes the row piece itself with 64- /
            // 128-bit stores -- no staging, no shared-memory round trip, a fifth of the store path's instructions (the
            // epilogue is bound by the ALU pipe: 2 warps per scheduler at one instruction per 2 cycles)
            char* gLane = reinterpret_cast<char*>(p.scores + (size_t)(ir.rowBase + (fG - f0) * kTileM + q * 32 + lane) * p.ld + mixBase);
            auto tiles = [&](auto spTag, auto dirTag) {
                constexpr int  SP     = decltype(spTag)::value;
                constexpr bool DIRECT = decltype(dirTag)::value && (MODE == 0 || NE == 16) && !BEST && SP > 0;
                constexpr int  R      = DIRECT ? SEGS / SP : 1;
                // VEC (R a multiple of 4): the R results go through the warp's staging block ([32 rows][R + 4] floats,
                // 128-bit writes) and leave as 128-bit stores of R / 4 lanes per row -- a store instruction then touches
                // 32 * 4 / R lines instead of 32 (the load/store unit takes one line per cycle: 512 -> 256 line slots per
                // tile at R = 8).  R = 2: the lane writes its 8 bytes itself.
                constexpr bool VEC    = DIRECT && R % 4 == 0;
                constexpr int  PV     = R + 4, WQ4 = R / 4 > 0 ? R / 4 : 1, RPP = 32 / WQ4;
                const uint32_t vqi = (uint32_t)lane % WQ4, vr0 = (uint32_t)lane / WQ4;
                const uint32_t vRead = stWarp + (vr0 * PV + vqi * 4) * 4, vWrite = stWarp + (uint32_t)lane * PV * 4;
                char* gVec = reinterpret_cast<char*>(p.scores + (size_t)(ir.rowBase + (fG - f0) * kTileM + q * 32 + vr0) * p.ld +
                                                     mixBase + vqi * 4);
                float          res[R];
                float* const   resp = DIRECT ? res : nullptr;
                float2 riNext = fG < f1 ? *rip : make_float2(0.f, 0.f);  // row term / scale, fetched one tile ahead
                for (uint32_t f = fG; f < f1; f += NG) {
                    const uint32_t t   = tItem + (f - f0);  // tile counter of the CTA (shadows the item-level one)
                    const uint32_t acc = t & 1;
                    const float    rt = riNext.x, rs = riNext.y;
                    rip += NG * kTileM;
                    if (f + NG < f1)
                        riNext = *rip;
                    if (p.dbg) ec0 = clock64();
                    mbar_wait(smem_u32(&done[t % kRing]), (t / kRing) & 1);
                    tc_fence_after();
                    if (p.dbg) { const long long c1 = clock64(); eW += c1 - ec0; ec0 = c1; }
                    const uint32_t taddr    = tmemBase + ((uint32_t)(q * 32) << 16) + acc * kTileN + part * PW;
                    const uint32_t drained  = smem_u32(&ready[(t + 2) % kRing]);  // tile t + 2 reuses this accumulator
                    EpiState st;
                    st.run = INFINITY, st.sum = 0.f, st.idx = 0xFFFFFFFFu, st.sp = stRow;
                    uint32_t va[32];
                    if (NE == 16) {
                        uint32_t vb[32];
                        // 64 columns: both loads in flight, then the accumulator goes straight back to the MMA warp
                        tmem_ld32(taddr, va);
                        tmem_ld32(taddr + 32, vb);
                        tmem_ld_wait();
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0)
                            mbar_arrive(drained);
                        if (MODE == 1 && SP > 0) {
                            auto get = [&](int i) { return __uint_as_float(i < 32 ? va[i & 31] : vb[i & 31]); };
                            if (sumFull)
                                epi_sum_uniform<BEST, (SP > 0 ? SP : 1), 8, true>(get, 0, part * PW, slots, liveBits, rt, rs, offB, st, resp);
                            else
                                epi_sum_uniform<BEST, (SP > 0 ? SP : 1), 8, false>(get, 0, part * PW, slots, liveBits, rt, rs, offB, st, resp);
                        }
                        else {
                            epi_chunk<MODE, BEST, SP>(va, 0, part * PW, slots, segEnd, segEmpty, liveBits, rt, rs, offB, st, resp);
                            epi_chunk<MODE, BEST, SP>(vb, 1, part * PW, slots, segEnd, segEmpty, liveBits, rt, rs, offB, st, resp);
                        }
                    }
                    else if (NG == 2) {
                        // 96 registers per thread (18 warps): one 32-column chunk live at a time; the scheduler's other
                        // warps (its second group among them) cover the TMEM load latency
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch) {
                            tmem_ld32(taddr + 32 * ch, va);
                            tmem_ld_wait();
                            if (ch == 3) {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0)
                                    mbar_arrive(drained);
                                if (p.dbg) { const long long c1 = clock64(); eL += c1 - ec0; ec0 = c1; }
                            }
                            epi_chunk<MODE, BEST, SP>(va, ch, part * PW, slots, segEnd, segEmpty, liveBits, rt, rs, offB, st, resp);
                        }
                    }
                    else {
                        // 128 columns in four 32-column chunks, the next load in flight while one is reduced
                        uint32_t vb[32];
                        tmem_ld32(taddr, va);
                        tmem_ld_wait();
                        tmem_ld32(taddr + 32, vb);
                        epi_chunk<MODE, BEST, SP>(va, 0, part * PW, slots, segEnd, segEmpty, liveBits, rt, rs, offB, st, resp);
                        tmem_ld_wait();
                        tmem_ld32(taddr + 64, va);
                        epi_chunk<MODE, BEST, SP>(vb, 1, part * PW, slots, segEnd, segEmpty, liveBits, rt, rs, offB, st, resp);
                        tmem_ld_wait();
                        tmem_ld32(taddr + 96, vb);
                        epi_chunk<MODE, BEST, SP>(va, 2, part * PW, slots, segEnd, segEmpty, liveBits, rt, rs, offB, st, resp);
                        tmem_ld_wait();
                        // accumulator drained into registers: hand it back to the MMA warp before finishing the tile
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0)
                            mbar_arrive(drained);
                        if (p.dbg) { const long long c1 = clock64(); eL += c1 - ec0; ec0 = c1; }
                        epi_chunk<MODE, BEST, SP>(vb, 3, part * PW, slots, segEnd, segEmpty, liveBits, rt, rs, offB, st, resp);
                    }
                    if (p.dbg) { const long long c1 = clock64(); eC += c1 - ec0; ec0 = c1; }
                    const uint32_t rbase = ir.rowBase + (f - f0) * kTileM + q * 32;
                    const uint32_t rows  = min(32u, ir.rowEnd > rbase ? ir.rowEnd - rbase : 0u);
                    if (VEC && p.directStores == 1) {
#pragma unroll
                        for (int i = 0; i < R / 4; ++i)
                            sts_v4(vWrite + i * 16, __float_as_uint(res[4 * i]), __float_as_uint(res[4 * i + 1]),
                                   __float_as_uint(res[4 * i + 2]), __float_as_uint(res[4 * i + 3]));
                        __syncwarp();
#pragma unroll
                        for (int ps = 0; ps < 32 / RPP; ++ps)
                            if (vr0 + ps * RPP < rows)
                                *reinterpret_cast<uint4*>(gVec + (size_t)ps * RPP * p.ld * 4) = lds_v4(vRead + ps * RPP * PV * 4);
                        gVec += gTileStep;
                        __syncwarp();
                        if (p.dbg) eS += clock64() - ec0;
                        continue;
                    }
                    if (DIRECT) {
                        if ((uint32_t)lane < rows) {
                            if (R >= 4) {
#pragma unroll
                                for (int i = 0; i < R / 4; ++i)
                                    reinterpret_cast<float4*>(gLane)[i] =
                                            make_float4(res[4 * i], res[4 * i + 1], res[4 * i + 2], res[4 * i + 3]);
                            }
                            else
                                *reinterpret_cast<float2*>(gLane) = make_float2(res[0], res[R > 1 ? 1 : 0]);
                        }
                        gLane += gTileStep;
                        if (p.dbg) eS += clock64() - ec0;
                        continue;
                    }
                    __syncwarp();
                    // coalesced store of the staged [32 rows x nMix] block: nMix passes of 32 consecutive elements;
                    // pointers advance by per-item constants (a pass that wraps past a row end takes the `w` step)
                    char*          gs    = gTile;
                    uint32_t       ss    = ssLane;
                    gTile += gTileStep;
                    if (adv_j == 0) {  // nMix divides 32: a lane keeps its column
    #pragma unroll 4
                        for (uint32_t i = 0, r = lr0; i < nMix; ++i, r += adv_r) {
                            if (r < rows) {
                                *reinterpret_cast<float*>(gs) = lds_f32(ss);
                                if (BEST)
                                    *reinterpret_cast<uint32_t*>(gs + bestDelta) = lds_u32(ss + offB);
                            }
                            gs += gStep, ss += sStep;
                        }
                    }
                    else {
                        uint32_t r = lr0, j = lj0;
                        for (uint32_t i = 0; i < nMix; ++i) {
                            if (r < rows) {
                                *reinterpret_cast<float*>(gs) = lds_f32(ss);
                                if (BEST)
                                    *reinterpret_cast<uint32_t*>(gs + bestDelta) = lds_u32(ss + offB);
                            }
                            const bool w = j + adv_j >= nMix;
                            j            = w ? j + adv_j - nMix : j + adv_j;
                            r += adv_r + (w ? 1u : 0u);
                            gs += w ? gStepW : gStep, ss += w ? sStepW : sStep;
                        }
                    }
                    __syncwarp();
                    if (p.dbg) eS += clock64() - ec0;
                }
            };
            // TileMeta::uni; the full-sum scorer is specialised on it with 16 epilogue warps (its default)
            const uint32_t uni = (MODE == 0 || NE == 16) ? ((meta[10] >> (8 * part)) & 0xFFu) : 0u;
            // row pieces of R = SEGS / uni floats: 16-byte stores for R >= 4, 8-byte for R = 2
            const uint32_t piece  = uni ? min(16u, (uint32_t)(SEGS / max(uni, 1u)) * 4u) : 0u;
            const bool     direct = (MODE == 0 || NE == 16) && !BEST && piece >= 8 && p.directStores &&
                                ((p.ld * 4) % piece) == 0 && ((mixBase * 4) % piece) == 0 &&
                                (reinterpret_cast<uintptr_t>(p.scores) % piece) == 0;
            auto run = [&](auto spTag) {
                // (full-sum scorer: 8 segments per part, so row pieces of at least two floats mean mixtures of <= 4 segments)
                if constexpr ((MODE == 0 || (NE == 16 && decltype(spTag)::value <= 4)) && !BEST && decltype(spTag)::value > 0) {
                    if (direct)
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
            // this warp no longer reads slot / meta of the B buffer
            if (lane == 0)
                mbar_arrive(smem_u32(&emptyB[b]));
        }
        if (p.dbg && lane == 0) {
            unsigned long long* d = p.dbg + blockIdx.x * 128;
            if (e == 0 && grp == 0)
                d[8] = eW, d[9] = eL, d[10] = eC, d[11] = eS;
            unsigned long long* w = d + 16 + (warp - EW0) * 4;  // every epilogue warp
            w[0] = eW, w[1] = eL, w[2] = eC, w[3] = eS;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmemBase));
    }
}

// 8 epilogue warps: 320 threads, up to 204 registers; 16 epilogue warps: 640 threads launched with 96 registers, the
// epilogue warps raise theirs to 112 (setmaxnreg, see EW0 in tc_score_body)
template<int SWZ, int MODE, bool BEST>
__global__ void __launch_bounds__(32 * (kEpiWarp0 + 8), 1) tc_score_kernel8(const KParams p) {
    tc_score_body<SWZ, MODE, BEST, 8>(p);
}
template<int SWZ, int MODE, bool BEST>
__global__ void __launch_bounds__(32 * ((BEST ? kEpiWarp0 : 4) + 16), 1) tc_score_kernel16(const KParams p) {  // see EW0
    tc_score_body<SWZ, MODE, BEST, 16>(p);
}
// two epilogue groups of eight warps (maximum approximation): 576 threads, 96 registers
template<int SWZ, int MODE, bool BEST>
__global__ void __launch_bounds__(32 * (kEpiWarp0 + 16), 1) tc_score_kernel8x2(const KParams p) {
    tc_score_body<SWZ, MODE, BEST, 8, 2>(p);
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
// LPF lanes share a frame (LPF = power of two >= the 8-element chunks per precision part, so 32 / LPF frames per
// warp): lane c of a frame's group owns chunk c of the hi and of the lo part; group reductions give the row term.
template<bool POOLED, int LPF>
__global__ void __launch_bounds__(256) tc_prep_kernel(const float* __restrict__ x, uint32_t T, uint32_t D,
                                                      uint32_t frameTiles, const float* __restrict__ ctr,
                                                      const float* __restrict__ scale, uint32_t kSteps, uint32_t atoms,
                                                      uint32_t swz, uint8_t* __restrict__ aTiles,
                                                      float2* __restrict__ rowInfo,
                                                      const uint32_t* __restrict__ rowMap, uint32_t fp8) {
    const uint32_t chunks = kSteps * 2;  // 8-element chunks per precision part
    const uint32_t epa    = swz / 2;     // fp16 elements per atom row
    const uint32_t lane   = threadIdx.x & 31, c = lane % LPF;
    const uint32_t kel    = POOLED ? D : 2 * D;
    constexpr uint32_t fpw = 32 / LPF;   // frames per warp
    const uint32_t framesPerGrid = gridDim.x * (blockDim.x >> 5) * fpw;
    const uint32_t rows = frameTiles * kTileM;
    for (uint32_t t = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * fpw + lane / LPF; t < rows;
         t += framesPerGrid) {
        float v[8];
        float vmax = 0.f, ssq = 0.f;
        // row t of the tile space holds frame `src` (subset scoring: tiles are aligned to segments, rowMap names the
        // frame of every row, 0xFFFFFFFF for padding)
        const uint32_t src = rowMap ? ((t % kTileM) < rowMap[2 * (t / kTileM) + 1] ? rowMap[2 * (t / kTileM)] + (t % kTileM) : 0xFFFFFFFFu)
                                    : (t < T ? t : 0xFFFFFFFFu);
        const float*   xr  = x + (size_t)(src == 0xFFFFFFFFu ? 0 : src) * D;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t k = c * 8 + j;
            v[j]             = 0.f;
            if (src != 0xFFFFFFFFu && c < chunks) {
                if (POOLED) {
                    if (k < D) {
                        v[j] = (xr[k] - ctr[k]) * scale[k];
                        ssq  = fmaf(v[j], v[j], ssq);
                    }
                }
                else if (k < D) {
                    const float y = (xr[k] - ctr[k]) * scale[k];
                    v[j]          = y * y;
                }
                else if (k < 2 * D) {
                    v[j] = (xr[k - D] - ctr[k - D]) * scale[k - D];
                }
            }
            if (k >= kel && k < kel + 3)
                v[j] = 1.f;  // multiplies the three fp16 pieces of the per-Gaussian constant (stored as 1/rowScale)
            vmax = fmaxf(vmax, fabsf(v[j]));
        }
#pragma unroll
        for (int o = LPF / 2; o > 0; o >>= 1) {
            vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
            ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
        }
        // power-of-two row scale: largest |a| / rs <= 16384 (non-finite frames are passed through clamped); with the
        // e4m3 correction operands <= 256, their largest finite value being 448
        float       rs    = 1.f;
        const float limit = fp8 ? 256.f : 16384.f;
        if (vmax > limit && vmax < 3.0e38f) {
            int e;
            frexpf(vmax, &e);  // vmax = m * 2^e, m in [0.5, 1)
            rs = ldexpf(1.f, e - (fp8 ? 8 : 14));
        }
        const float inv = 1.f / rs;
        if (c == 0)
            rowInfo[t] = make_float2(ssq, rs);
        const uint32_t tile = t / kTileM, row = t % kTileM;
        uint8_t*       base = aTiles + (size_t)tile * atoms * kTileM * swz;
        if (c < chunks) {
            __align__(16) __half hi[8], lo[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                split_fp16(v[j] * inv, hi[j], lo[j]);
            const uint32_t kHi = c * 8, kLo = kSteps * 16 + c * 8;
            *reinterpret_cast<uint4*>(base + (size_t)(kHi / epa) * kTileM * swz +
                                      swz_offset(row, (kHi % epa) / 8, swz)) = *reinterpret_cast<const uint4*>(hi);
            if (fp8) {
                // part 2, k-step c / 2: 16-byte chunk of the lo bytes, then the chunk of the hi bytes; this lane's 8
                // dimensions are the lower or the upper half of both
                __align__(8) uint8_t l8[8], h8[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float a = v[j] * inv;
                    l8[j] = (uint8_t)__nv_cvt_float_to_fp8((a - __half2float(hi[j])) * kLoScale, __NV_SATFINITE, __NV_E4M3);
                    h8[j] = (uint8_t)__nv_cvt_float_to_fp8(a, __NV_SATFINITE, __NV_E4M3);
                }
                const uint32_t cpa = swz / 16, cLo = kSteps * 2 + (c / 2) * 2, cHi = cLo + 1, half8 = (c & 1) * 8;
                *reinterpret_cast<uint2*>(base + (size_t)(cLo / cpa) * kTileM * swz + swz_offset(row, cLo % cpa, swz) + half8) =
                        *reinterpret_cast<const uint2*>(l8);
                *reinterpret_cast<uint2*>(base + (size_t)(cHi / cpa) * kTileM * swz + swz_offset(row, cHi % cpa, swz) + half8) =
                        *reinterpret_cast<const uint2*>(h8);
            }
            else
                *reinterpret_cast<uint4*>(base + (size_t)(kLo / epa) * kTileM * swz +
                                          swz_offset(row, (kLo % epa) / 8, swz)) = *reinterpret_cast<const uint4*>(lo);
        }
        // zero the unused tail of the last atom
        const uint32_t used = chunks * 2, all = atoms * (swz / 16);  // 16-byte chunks per row
        for (uint32_t cc = used + c; cc < all; cc += LPF)
            *reinterpret_cast<uint4*>(base + (size_t)(cc / (swz / 16)) * kTileM * swz +
                                      swz_offset(row, cc % (swz / 16), swz)) = make_uint4(0, 0, 0, 0);
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

// NE = 8 / 16: one epilogue group of that many warps; NE = 82: two groups of eight
template<int SWZ, int MODE, bool BEST, int NE>
cudaError_t launch_tc(const KParams& p, int grid, size_t smemBytes, cudaStream_t stream) {
    void (*kernel)(const KParams) = nullptr;
    if constexpr (NE == 16)
        kernel = tc_score_kernel16<SWZ, MODE, BEST>;
    else if constexpr (NE == 82 && MODE == 0)
        kernel = tc_score_kernel8x2<SWZ, MODE, BEST>;
    else
        kernel = tc_score_kernel8<SWZ, MODE, BEST>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);
    if (e != cudaSuccess)
        return e;
    kernel<<<grid, (NE == 16 && !BEST) ? 32 * (4 + 16) : 32 * (kEpiWarp0 + ((NE == 8 || (NE == 82 && MODE != 0)) ? 8 : 16)),
             smemBytes, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace

// ---- host side: packing --------------------------------------------------------------------------------
static uint32_t smemNeed(uint32_t bTileBytes, uint32_t aTileBytes, uint32_t buffersB, uint32_t stagesA,
                         uint32_t pitch, bool best, uint32_t epiWarps) {
    const uint32_t bBytes = (bTileBytes + 1023u) & ~1023u;
    return 1024 + buffersB * bBytes + stagesA * aTileBytes + epiWarps * 32 * pitch * 4 * (best ? 2 : 1) + 512;
}

int gmm_tc_pack(gmm_handle* h, const uint32_t* offsets, const float* mean, const float* isr, const uint32_t* covIdx,
                const float* w2, const float* logNorm, bool force) {
    TensorModel& tm = h->tensor;
    tm              = TensorModel();
    const uint32_t M = h->M, G = h->G, D = h->D, C = h->C;
    if (G == 0 || M == 0)
        return GMM_OK;
    // 64-byte swizzle atoms by default: an operand row of 45 dimensions is 192 bytes = three 64-byte atoms (two 128-byte
    // atoms would carry 64 bytes of padding per row), which leaves room for four A stages beside two B buffers instead of
    // two -- 4501x16: 0.391 -> 0.368 ms per 32000 frames, 12001x64: 1.086 -> 1.033 ms.  GMM_TC_SWIZZLE=128 selects the other.
    const char* env = getenv("GMM_TC_SWIZZLE");
    tm.swz          = (env && atoi(env) == 128) ? 128 : 64;
    tm.pooled       = (C == 1);
    bool fp8Forced  = false;
    {
        // Default: fp16 hi.hi + ONE e4m3 product for both corrections wherever the guard below allows it (pooled models);
        // "fp16": always three fp16 products; "fp8!": the e4m3 split regardless of the guard (experiments)
        const char* split = getenv("GMM_TC_SPLIT");
        fp8Forced         = split && !strcmp(split, "fp8!");
        tm.fp8            = !(split && !strcmp(split, "fp16")) && tm.pooled;
    }
    const uint32_t kel = tm.pooled ? D : 2 * D;
    tm.kSteps          = (kel + 3 + 15) / 16;  // 3 spare K columns carry the per-Gaussian constant
    if (tm.kSteps > (uint32_t)kMaxKSteps)
        return GMM_OK;  // dimension beyond the tensor kernel: tm.valid stays false
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
    std::vector<uint32_t> mixPerPart;
    auto isLive = [&](uint32_t g) { return isfinite(w2[g]) && isfinite(logNorm[covIdx[g]]); };
    // column part = the columns one epilogue warp reduces: 64 (16 epilogue warps) when every mixture fits, else 128
    uint32_t maxLive = 0;
    for (uint32_t m = 0; m < M; ++m) {
        uint32_t n = 0;
        for (uint32_t g = offsets[m]; g < offsets[m + 1]; ++g)
            n += isLive(g);
        maxLive = std::max(maxLive, n);
    }
    const char* envW = getenv("GMM_TC_EPI_WARPS");
    // 16 warps pay off where the epilogue is latency-bound (log-sum-exp); for the min tree their extra issue
    // slots starve the MMA-issuing warp that shares a scheduler with a quarter of them (profiles/ubench_mma.cu)
    const bool wide  = envW ? atoi(envW) == 16 : h->mode == GMM_MODE_SUM;
    tm.partW         = (maxLive <= 64 && wide) ? 64 : 128;
    const uint32_t partW = tm.partW, partsPerTile = kTileN / partW;
    auto partOf = [&]() { return (uint32_t)(cols.size() / partW); };
    auto ensurePart = [&](uint32_t pt) {
        while (metas.size() * partsPerTile <= pt) {
            TileMeta z = {0, 0, {0, 0, 0, 0}, {0, 0, 0, 0}, 0};
            metas.push_back(z);
        }
        while (mixPerPart.size() <= pt)
            mixPerPart.push_back(0);
    };
    for (uint32_t m = 0; m < M; ++m) {
        std::vector<uint32_t> live;
        for (uint32_t g = offsets[m]; g < offsets[m + 1]; ++g)
            if (isLive(g))
                live.push_back(g);
        uint32_t padded = std::max<uint32_t>(8, (uint32_t)((live.size() + 7) / 8 * 8));
        if (padded > 128 || offsets[m + 1] - offsets[m] > 0xFFFF)
            return GMM_OK;  // does not fit the tensor kernel: tm.valid stays false
        const uint32_t inPart = (uint32_t)(cols.size() % partW);
        if (inPart + padded > partW) {
            const Col pad = {0xFFFFFFFFu, 0xFFFF};
            cols.insert(cols.end(), partW - inPart, pad);
        }
        const uint32_t pt = partOf();
        ensurePart(pt);
        TileMeta& mt = metas[pt / partsPerTile];
        if (mixPerPart[pt] == 0)
            mt.mixBase[pt % partsPerTile] = m;
        mixPerPart[pt]++;
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
    ensurePart(tm.nTiles * partsPerTile - 1);
    // uniform parts: all segments of the part taken by mixtures of the same power-of-two segment count, none empty
    for (TileMeta& mt : metas) {
        const uint32_t segsPerPart = partW / 8;
        for (uint32_t pp = 0; pp < partsPerTile; ++pp) {
            const uint32_t mask = segsPerPart == 32 ? 0xFFFFFFFFu : ((1u << segsPerPart) - 1u);
            const uint32_t ends = (mt.segEnd >> (pp * segsPerPart)) & mask, empt = (mt.segEmpty >> (pp * segsPerPart)) & mask;
            for (uint32_t S = 1; S <= 8 && S <= segsPerPart; S *= 2) {
                uint32_t pat = 0;
                for (uint32_t b = S - 1; b < segsPerPart; b += S)
                    pat |= 1u << b;
                if (ends == pat && empt == 0)
                    mt.uni |= S << (8 * pp);
            }
        }
    }
    tm.maxMixPerPart = 1;
    for (uint32_t v : mixPerPart)
        tm.maxMixPerPart = std::max(tm.maxMixPerPart, v);
    {   // regular model (see TensorModel): same live count everywhere, padded width dividing the column part
        const uint32_t n0 = offsets[1] - offsets[0], P = std::max<uint32_t>(8, (n0 + 7) / 8 * 8);
        bool           reg = n0 > 0 && partW % P == 0 && M >= kTileN / P;
        for (uint32_t m = 0; m < M && reg; ++m) {
            reg = offsets[m + 1] - offsets[m] == n0;
            for (uint32_t g = offsets[m]; g < offsets[m + 1] && reg; ++g)
                reg = isLive(g);
        }
        tm.regular    = reg;
        tm.mixCols    = P;
        tm.mixPerTile = kTileN / P;
    }

    // ---- shared-memory configuration (sized for the arg-min variant) ----
    const uint32_t pitch = (tm.maxMixPerPart + 3) / 4 * 4 + 5, epiWarps = 4 * partsPerTile;  // odd, >= R + 4 (vector stores)
    tm.buffersB = 0;
    for (uint32_t nb = 2; nb >= 1 && !tm.buffersB; --nb)
        for (uint32_t sa = kMaxStagesA; sa >= 2; --sa)
            if (smemNeed(tm.tileBytes, tm.aTileBytes, nb, sa, pitch, true, epiWarps) <= kSmemLimit) {
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
    if (tm.fp8 && !fp8Forced) {
        // The e4m3 corrections hold the 1e-4 bar only for compact models.  Predicted error of a score:
        // the correction products carry relative rounding errors on terms of size 2^-12 |a_d b_d|, about
        // 2^-13 sqrt(sum_d a_d^2 b_d^2) per score when the roundings are independent; for a frame drawn near Gaussian b
        // (a = b + unit noise, the pairs with the SMALLEST scores) that is 2^-13 q(b), q(b)^2 = sum_d (b_d^2 + 1) b_d^2,
        // against a score of at least sMin = (min_g (w2 + L) + D - 3 sqrt(2D)) / 2.  Emulation over D = 9..64, mean spreads
        // 0.5..4 sigma and 21 seeds at the 2-sigma spread of the benchmark models (profiles/split_guard_r02.txt) finds the
        // largest relative error of a whole batch at <= 0.33 of  pred = 2^-13 rms_g q / sMin  for D >= 16 (<= 0.51 at D = 9):
        // the split is taken for pred <= 1.5e-4 (D >= 16; expected worst error 5e-5) or 1.0e-4 (9 <= D < 16); cfg1 / cfg3:
        // pred 1.3e-4, emulated errors 1.9..3.7e-5.  D < 9 averages too few roundings: three products.
        // (e5m2 operands, the first version of this split, reached 1.13e-4 on one of those seeds: not enough margin.)
        double q2 = 0.0, cMin = 1e300;
        size_t n  = 0;
        for (const Col& c : cols)
            if (c.g != 0xFFFFFFFFu && c.slot != 0xFFFF) {
                double q = 0.0;
                for (uint32_t d = 0; d < D; ++d) {
                    const double b = ((double)mean[(size_t)c.g * D + d] - (double)ctrF[d]) * (double)isr[d];
                    q += (b * b + 1.0) * b * b;
                }
                q2 += q;
                cMin = std::min(cMin, (double)w2[c.g] + (double)logNorm[covIdx[c.g]]);
                ++n;
            }
        const double sMin = std::max(1.0, 0.5 * cMin + 0.5 * std::max(0.0, (double)D - 3.0 * sqrt(2.0 * D)));
        const double pred = n ? ldexp(sqrt(q2 / (double)n), -13) / sMin : 0.0;
        if (D < 9 || !(pred <= (D >= 16 ? 1.5e-4 : 1.0e-4)))
            tm.fp8 = false;
    }
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
                if (tm.fp8) {
                    // part 2, k-step k / 16: chunk of the hi bytes, then chunk of the lo bytes (mirrors tc_prep_kernel)
                    __half hh;
                    memcpy(&hh, &hi, 2);
                    const float    v   = std::min(std::max(row[k], -60000.f), 60000.f);
                    const uint8_t  h8  = (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3);
                    const uint8_t  l8  = (uint8_t)__nv_cvt_float_to_fp8((v - __half2float(hh)) * kLoScale, __NV_SATFINITE, __NV_E4M3);
                    const uint32_t cpa = tm.swz / 16, cH = tm.kSteps * 2 + (k / 16) * 2, cL = cH + 1;
                    base[(size_t)(cH / cpa) * kTileN * tm.swz + swz_offset(r, cH % cpa, tm.swz) + k % 16] = h8;
                    base[(size_t)(cL / cpa) * kTileN * tm.swz + swz_offset(r, cL % cpa, tm.swz) + k % 16] = l8;
                }
                else
                    memcpy(pLo, &lo, 2);
            }
        }
    }
    // The expanded form |a|^2 + |b|^2 - 2 a.b cancels in f32: its absolute error grows like 2^-22 * |b|^2 while a
    // frame next to such a mean has a score of order D/2.  Models whose centred, whitened means are further out
    // than 8 sigma (rms per dimension) keep the 1e-4 relative bar only in the direct kernel.
    tm.accuracyOk = !(kappa > 64.0 * D);  // reported even when the tensor kernel is forced (gmm_float_accuracy_ok)
    if (!force && !tm.accuracyOk)
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

namespace {

// One run of the preparation + scoring kernels.  Full scoring: FT = ceil(T / 128) frame tiles over all B tiles of the
// model.  Subset scoring: tile-aligned row space described by rowMap, items from itemTab, B tiles gathered per call.
struct TcJob {
    uint32_t        T, FT;
    const float*    d_x;
    const uint32_t* rowMap;   // device, per frame tile {first frame, frames}, or null
    const uint8_t*  bTiles;
    uint32_t        nTiles;
    const uint32_t* itemTab;  // device, 8 words per item, or null
    uint32_t        nItems;
    float*          scores;
    uint32_t*       best;
    size_t          ld;
    bool            prep = true;  // false: the frame images of the previous run over the same frames are reused
    uint32_t        prepT = 0;    // prep: frames to prepare from d_x on (0: T); the run itself may cover fewer
    uint32_t        aTile0 = 0;   // first frame tile of the run among the prepared ones
};

int tc_run(gmm_handle* h, const TcJob& job) {
    const TensorModel& tm = h->tensor;
    const uint32_t     FT = job.FT, T = job.T;
    const float*       d_x = job.d_x;
    float*             d_scores = job.scores;
    uint32_t*          d_best   = job.best;
    const size_t       ld       = job.ld;
    void *         aT = nullptr, *rT = nullptr;
    int            rc;
    const uint32_t prepT  = job.prepT ? job.prepT : T;
    const uint32_t prepFT = job.prepT ? (job.prepT + kTileM - 1) / kTileM : FT;
    const uint32_t holdFT = std::max(job.aTile0 + FT, job.prep ? prepFT : 0u);
    if ((rc = h->reserve(SCR_A, (size_t)holdFT * tm.aTileBytes, &aT)))
        return rc;
    if ((rc = h->reserve(SCR_ROW, (size_t)holdFT * kTileM * 8, &rT)))
        return rc;
    if (job.prep) {
        const uint32_t T = prepT, FT = prepFT;  // the preparation may run ahead of the frames scored below
        const uint32_t chunks = tm.kSteps * 2;
        const uint32_t lpf    = chunks <= 8 ? 8 : chunks <= 16 ? 16 : 32;
        const uint32_t fpb    = 8 * (32 / lpf);  // frames per 256-thread block
        int blocks = (int)std::min<uint32_t>((FT * kTileM + fpb - 1) / fpb, (uint32_t)h->smCount * 16);
#define GMM_PREP(P, L)                                                                                          \
    tc_prep_kernel<P, L><<<blocks, 256, 0, h->stream>>>(d_x, T, h->D, FT, tm.center, tm.scale, tm.kSteps, tm.nAtoms, \
                                                        tm.swz, (uint8_t*)aT, (float2*)rT, job.rowMap, tm.fp8 ? 1u : 0u)
        if (tm.pooled) {
            if (lpf == 8)
                GMM_PREP(true, 8);
            else if (lpf == 16)
                GMM_PREP(true, 16);
            else
                GMM_PREP(true, 32);
        }
        else {
            if (lpf == 8)
                GMM_PREP(false, 8);
            else if (lpf == 16)
                GMM_PREP(false, 16);
            else
                GMM_PREP(false, 32);
        }
#undef GMM_PREP
        h->launches++;
        GMM_CUDA_TRY(h, cudaGetLastError());
    }
    KParams p;
    p.bTiles     = job.bTiles;
    p.itemTab    = job.itemTab;
    p.nItems     = job.nItems;
    p.aTiles     = (const uint8_t*)aT + (size_t)job.aTile0 * tm.aTileBytes;
    p.rowInfo    = (const float2*)rT + (size_t)job.aTile0 * kTileM;
    p.scores     = d_scores;
    p.best       = d_best;
    p.ld         = ld;
    p.T          = T;
    p.nTiles     = job.nTiles;
    p.frameTiles = FT;
    p.kSteps     = tm.kSteps;
    p.fp8        = tm.fp8 ? 1u : 0u;
    p.atoms      = tm.nAtoms;
    p.aTileBytes = tm.aTileBytes;
    p.bTileBytes = tm.tileBytes;
    p.buffersB   = tm.buffersB;
    p.stagePitch = (tm.maxMixPerPart + 3) / 4 * 4 + 5;
    p.dbg        = nullptr;
    {
        // row pieces of uniform tiles: 1 = 128-bit stores through the staging block (default), 2 = every lane writes its own
        // piece (GMM_TC_STORES=lane), 0 = the general staged scalar path (GMM_TC_STORES=staged)
        const char* e  = getenv("GMM_TC_STORES");
        p.directStores = (e && !strcmp(e, "staged")) ? 0u : (e && !strcmp(e, "lane")) ? 2u : 1u;
        const char* sf = getenv("GMM_TC_SUM_FULL");
        p.sumFull      = (sf && sf[0] == '0') ? 0u : 1u;
    }
    const bool debug = getenv("GMM_TC_DEBUG") != nullptr;
    if (debug) {
        void* d = nullptr;
        if ((rc = h->reserve(SCR_MISC, (size_t)h->smCount * 1024, &d)))
            return rc;
        p.dbg = (unsigned long long*)d;
    }
    const uint32_t epiWarps = 4 * (kTileN / tm.partW);
    const bool best = d_best != nullptr;
    // GMM_TC_GROUPS=2: two eight-warp epilogue groups on alternate tiles.  Measured and left off: with the reduction of one
    // group always running beside the MMAs the tensor instructions retire slower (issue + commit 694 -> 995 cycles per
    // tile on the 4501x16 model; 0.385 -> 0.550 ms per 32000 frames), although the same arrangement helps the quantized
    // kernel, whose tile is not bound by the tensor pipe (gmm_qtc.cu)
    bool twoGroups = false;
    if (const char* e = getenv("GMM_TC_GROUPS"))
        twoGroups = epiWarps == 8 && h->mode != GMM_MODE_SUM && atoi(e) == 2;
    const uint32_t stageWarps = twoGroups ? 16 : epiWarps;
    // A ring depth: whatever shared memory allows for this variant
    p.stagesA = tm.stagesA;
    while (p.stagesA < (uint32_t)kMaxStagesA &&
           smemNeed(tm.tileBytes, tm.aTileBytes, tm.buffersB, p.stagesA + 1, p.stagePitch, best, stageWarps) <= kSmemLimit)
        p.stagesA++;
    while (p.stagesA > 2 &&
           smemNeed(tm.tileBytes, tm.aTileBytes, tm.buffersB, p.stagesA, p.stagePitch, best, stageWarps) > kSmemLimit)
        p.stagesA--;  // the second group's staging blocks may cost an A stage
    if (twoGroups && smemNeed(tm.tileBytes, tm.aTileBytes, tm.buffersB, p.stagesA, p.stagePitch, best, stageWarps) > kSmemLimit)
        twoGroups = false;
    const size_t smemBytes = smemNeed(tm.tileBytes, tm.aTileBytes, tm.buffersB, p.stagesA, p.stagePitch, best,
                                      twoGroups ? 16 : epiWarps);
    // split the frame range so that the items fill the SMs evenly: cost in A-tile MMA times.  Items are handed out
    // round-robin, so the CTAs running together hold ADJACENT B tiles over the same frames and their 16..64-byte row
    // pieces merge into full lines in L2.  (Giving every CTA one contiguous range of the linearised tile space
    // balances perfectly but scatters those pieces in time: measured 0.49 -> 0.82 ms on the 4501x16 model.)
    const double bCost = tm.buffersB == 2 ? 0.5 : 2.0;
    uint32_t     bestS = 1;
    double       bestCost = 1e300;
    for (uint32_t S = 1; S <= std::min<uint32_t>(FT, 64); ++S) {
        const uint32_t per = (FT + S - 1) / S;
        if ((uint64_t)per * (S - 1) >= FT)
            continue;  // an empty chunk
        const uint64_t items  = (uint64_t)job.nTiles * S;
        const uint64_t rounds = (items + h->smCount - 1) / h->smCount;
        const double   cost   = (double)rounds * (per + bCost);
        if (cost < bestCost - 1e-9) {
            bestCost = cost;
            bestS    = S;
        }
    }
    p.split        = bestS;
    const int grid = (int)std::min<uint64_t>((uint64_t)h->smCount,
                                             job.itemTab ? (uint64_t)job.nItems : (uint64_t)job.nTiles * bestS);
    // The model's last B tile is usually a partial one (pad columns: the ragged, slower epilogue path).  With whole frame
    // ranges per item (split 1) it would be the LAST item of a CTA that already has the most items; it trades places
    // with a first-round tile of a CTA that has one item fewer.
    p.swapTile = job.nTiles - 1;
    if (!job.itemTab && bestS == 1 && job.nTiles > (uint32_t)grid && job.nTiles % grid != 0)
        p.swapTile = grid - 1;
    cudaError_t e;
    const int   mode = h->mode == GMM_MODE_SUM ? 1 : 0;
#define GMM_TC_LAUNCH(SWZ, NE)                                                                               \
    (mode == 0 ? (best ? launch_tc<SWZ, 0, true, NE>(p, grid, smemBytes, h->stream)                           \
                       : launch_tc<SWZ, 0, false, NE>(p, grid, smemBytes, h->stream))                         \
               : (best ? launch_tc<SWZ, 1, true, NE>(p, grid, smemBytes, h->stream)                           \
                       : launch_tc<SWZ, 1, false, NE>(p, grid, smemBytes, h->stream)))
    h->timeBegin();
    if (tm.swz == 128)
        e = epiWarps == 16 ? GMM_TC_LAUNCH(128, 16) : twoGroups ? GMM_TC_LAUNCH(128, 82) : GMM_TC_LAUNCH(128, 8);
    else
        e = epiWarps == 16 ? GMM_TC_LAUNCH(64, 16) : twoGroups ? GMM_TC_LAUNCH(64, 82) : GMM_TC_LAUNCH(64, 8);
#undef GMM_TC_LAUNCH
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, e);
    if (debug) {  // development aid: where the MMA-issuing thread spends its cycles
        std::vector<unsigned long long> d((size_t)h->smCount * 128);
        GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
        GMM_CUDA_TRY(h, cudaMemcpy(d.data(), p.dbg, d.size() * 8, cudaMemcpyDeviceToHost));
        double tot = 0, wR = 0, wI = 0, wC = 0, tiles = 0;
        double eW = 0, eL = 0, eC = 0, eS = 0;
        for (int c = 0; c < grid; ++c) {
            tot += d[c * 128], wR += d[c * 128 + 1], wI += d[c * 128 + 2], wC += d[c * 128 + 3], tiles += d[c * 128 + 4];
            eW += d[c * 128 + 8], eL += d[c * 128 + 9], eC += d[c * 128 + 10], eS += d[c * 128 + 11];
        }
        tiles = std::max(1.0, tiles);
        {
            unsigned long long g0 = ~0ull, g1 = 0;
            double dmin = 1e30, dmax = 0, dsum = 0, ns = 0;
            for (int c = 0; c < grid; ++c) {
                g0 = std::min(g0, d[c * 128 + 6]), g1 = std::max(g1, d[c * 128 + 7]);
                const double dur = (double)(d[c * 128 + 7] - d[c * 128 + 6]) * 1e-3;
                dmin = std::min(dmin, dur), dmax = std::max(dmax, dur), dsum += dur, ns += dur * 1e3;
            }
            fprintf(stderr, "[gmm_tc] MMA loops: %.1f us first start to last end; per CTA %.1f / %.1f / %.1f us (min / mean / max); "
                    "SM clock %.0f MHz (cycles / globaltimer)\n",
                    (double)(g1 - g0) * 1e-3, dmin, dsum / grid, dmax, ns > 0 ? tot / ns * 1e3 : 0.0);
        }
        for (uint32_t w = 0; w < epiWarps * (twoGroups ? 2 : 1); ++w) {
            double v[4] = {0, 0, 0, 0};
            for (int c = 0; c < grid; ++c)
                for (int k = 0; k < 4; ++k)
                    v[k] += d[c * 128 + 16 + w * 4 + k];
            fprintf(stderr, "[gmm_tc]   epilogue warp %2u: wait-for-MMA %.0f + drain %.0f + last chunk %.0f + stores %.0f = %.0f\n", w,
                    v[0] / tiles, v[1] / tiles, v[2] / tiles, v[3] / tiles, (v[0] + v[1] + v[2] + v[3]) / tiles);
        }
        fprintf(stderr,
                "[gmm_tc] epilogue warp 0, cycles per tile: wait-for-MMA %.0f + TMEM drain (loads, first three chunks) %.0f + "
                "last chunk %.0f + stores %.0f\n",
                eW / tiles, eL / tiles, eC / tiles, eS / tiles);
        fprintf(stderr,
                "[gmm_tc] grid %d split %u stagesA %u buffersB %u kSteps %u epiWarps %u x %d | MMA thread, cycles per tile: "
                "total %.0f = wait-ready %.0f + issue+commit %.0f + wait for the B tile %.0f\n",
                grid, p.split, p.stagesA, p.buffersB, p.kSteps, epiWarps, twoGroups ? 2 : 1, tot / tiles, wR / tiles, wI / tiles,
                wC / tiles);
    }
    return GMM_OK;
}

}  // namespace

// B tiles [tile0, tile1) of the model over one launch worth of frames (gmm_gather.cu scores a shard one column range
// at a time); prep = false reuses the frame images an earlier call left in the scratch.  A call over a range of
// frames passes the range's frames as d_x / T / d_scores and names its first frame tile among the prepared ones
// (frameTile0); the preparing call passes the start of ALL frames as d_x and their number as prepFrames.
int gmm_tc_score_tiles(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, size_t ld, uint32_t tile0,
                       uint32_t tile1, bool prep, uint32_t prepFrames, uint32_t frameTile0) {
    if (T == 0 || tile1 <= tile0)
        return GMM_OK;
    TcJob job;
    job.T       = T;
    job.FT      = (T + kTileM - 1) / kTileM;
    job.d_x     = d_x;
    job.rowMap  = nullptr;
    job.bTiles  = h->tensor.bTiles + (size_t)tile0 * h->tensor.tileBytes;
    job.nTiles  = tile1 - tile0;
    job.itemTab = nullptr;
    job.nItems  = 0;
    job.scores  = d_scores;
    job.best    = nullptr;
    job.ld      = ld;
    job.prep    = prep;
    job.prepT   = prepFrames;
    job.aTile0  = frameTile0;
    return tc_run(h, job);
}

int gmm_tc_score(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld) {
    // Long batches go out as launches of at most kChunk frames: the A tiles of one launch (256 bytes per frame) and
    // the model's B tiles then stay in L2 across the B-tile passes, instead of being streamed from HBM once per pass
    // (4501x16 model, sustained: 57.0 M frames/s in one 512000-frame launch, 62.3 M in 32000-frame launches).
    static const uint32_t kChunk = [] {
        const char* e = getenv("GMM_TC_CHUNK");
        const long  v = e ? atol(e) : 32768;
        return (uint32_t)std::max<long>(kTileM, v / kTileM * kTileM);
    }();
    for (uint32_t t0 = 0; t0 < T; t0 += kChunk) {
        const uint32_t n = std::min(kChunk, T - t0);
        TcJob          job;
        job.T       = n;
        job.FT      = (n + kTileM - 1) / kTileM;
        job.d_x     = d_x + (size_t)t0 * h->D;
        job.rowMap  = nullptr;
        job.bTiles  = h->tensor.bTiles;
        job.nTiles  = h->tensor.nTiles;
        job.itemTab = nullptr;
        job.nItems  = 0;
        job.scores  = d_scores + (size_t)t0 * ld;
        job.best    = d_best ? d_best + (size_t)t0 * ld : nullptr;
        job.ld      = ld;
        const int rc = tc_run(h, job);
        if (rc)
            return rc;
    }
    return GMM_OK;
}

// ---- subset scoring --------------------------------------------------------------------------------------------
// The reference's context scorers evaluate a mixture only when a consumer asks for it (CachedContextScorer,
// /root/reference/src/Mm/FeatureScorer.hh:170-230): an aligner touches the few hundred emissions of the segment's
// automaton, not all M.  Here every segment gets its own B tiles, assembled from the canonical image one mixture
// (mixCols rows = whole swizzle atoms) at a time, and an item table pairs them with the segment's frame tiles; the
// scoring kernel itself is the one of the full path, so a (frame, mixture) score is bit-identical in both.
namespace {

__global__ void __launch_bounds__(256) gather_tiles_kernel(const uint8_t* __restrict__ bTiles, uint32_t tileBytes,
                                                           uint32_t atoms, uint32_t swz, uint32_t mixCols,
                                                           uint32_t mixPerTile, uint32_t mixPerPart, uint32_t parts,
                                                           const uint32_t* __restrict__ vtMix,    // [nVt x mixPerTile]
                                                           const uint32_t* __restrict__ vtBase,   // [nVt] first column
                                                           uint32_t nVt, uint8_t* __restrict__ out) {
    const uint32_t c16PerRow = swz / 16, operandU4 = atoms * kTileN * c16PerRow;
    for (uint32_t vt = blockIdx.x; vt < nVt; vt += gridDim.x) {
        uint8_t* dst = out + (size_t)vt * tileBytes;
        for (uint32_t i = threadIdx.x; i < operandU4; i += blockDim.x) {
            const uint32_t a = i / (kTileN * c16PerRow), rem = i % (kTileN * c16PerRow);
            const uint32_t row = rem / c16PerRow, c = rem % c16PerRow;
            const uint32_t m    = vtMix[vt * mixPerTile + row / mixCols];
            const uint32_t srow = (m % mixPerTile) * mixCols + row % mixCols;
            // rows move by multiples of 8 (mixCols is one), so the 16-byte chunk keeps its swizzled place in the row
            const uint4* src = reinterpret_cast<const uint4*>(bTiles + (size_t)(m / mixPerTile) * tileBytes +
                                                              (size_t)a * kTileN * swz + (size_t)srow * swz) + c;
            reinterpret_cast<uint4*>(dst + (size_t)a * kTileN * swz + (size_t)row * swz)[c] = __ldg(src);
        }
        // slot table and TileMeta of a full canonical tile; the mixture base of every column part is this tile's
        const uint32_t  metaOff = atoms * kTileN * swz;
        const uint32_t* msrc    = reinterpret_cast<const uint32_t*>(bTiles + metaOff);
        uint32_t*       mdst    = reinterpret_cast<uint32_t*>(dst + metaOff);
        for (uint32_t i = threadIdx.x; i < kMetaBytes / 4; i += blockDim.x) {
            uint32_t       v = msrc[i];
            const uint32_t w = i - kTileN * 2 / 4;  // word index inside TileMeta (slots come first)
            if (i >= kTileN * 2 / 4 && w >= 2 && w < 2 + parts)
                v = vtBase[vt] + (w - 2) * mixPerPart;
            mdst[i] = v;
        }
    }
}

}  // namespace

bool gmm_tc_subset_capable(const gmm_handle* h) {
    return h->floatValid && h->tensor.valid && h->tensor.regular && h->kernel == GMM_KERNEL_TENSOR;
}

// leading dimension a compact score block needs for segments whose longest emission list has maxListLength entries
uint32_t gmm_tc_subset_ld(const gmm_handle* h, uint32_t maxListLength) {
    const uint32_t mpt = std::max(1u, h->tensor.mixPerTile);
    return std::max(1u, (maxListLength + mpt - 1) / mpt) * mpt;
}

int gmm_tc_score_subset(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const uint32_t* emOff,
                        const uint32_t* emList, const float* d_x, float* d_scores, size_t ld) {
    const TensorModel& tm = h->tensor;
    if (!gmm_tc_subset_capable(h))
        return GMM_ERR_UNSUPPORTED;
    const uint32_t mpt = tm.mixPerTile, parts = kTileN / tm.partW, mixPerPart = tm.partW / tm.mixCols;
    // tile-aligned row space, virtual tiles and items, segment by segment; the tables are built in one pinned block
    // laid out like its device copy (one asynchronous copy per call)
    uint32_t FT = 0, nVt = 0;
    for (uint32_t s = 0; s < nSeg; ++s) {
        const uint32_t f0 = frameOffsets[s], f1 = frameOffsets[s + 1], n = emOff[s + 1] - emOff[s];
        if (f1 <= f0 || n == 0)
            continue;
        if ((size_t)gmm_tc_subset_ld(h, n) > ld) {
            h->setError("subset scoring: score block narrower than the segment's emission list");
            return GMM_ERR_ARG;
        }
        FT += (f1 - f0 + kTileM - 1) / kTileM;
        nVt += (n + mpt - 1) / mpt;
    }
    if (nVt == 0)
        return GMM_OK;
    auto         al   = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t oMap = 0, oMix = al((size_t)FT * 8), oBase = oMix + al((size_t)nVt * mpt * 4),
                 oItem = oBase + al((size_t)nVt * 4), total = oItem + al((size_t)nVt * 32);
    void *tab = nullptr, *vtiles = nullptr, *pin = nullptr;
    int   rc;
    if ((rc = h->reservePinned(PIN_SUBSET, total, &pin)))
        return rc;
    uint32_t *tileMap = (uint32_t*)((char*)pin + oMap), *vtMix = (uint32_t*)((char*)pin + oMix),
             *vtBase = (uint32_t*)((char*)pin + oBase), *items = (uint32_t*)((char*)pin + oItem);
    uint32_t ft = 0, vt = 0;
    for (uint32_t s = 0; s < nSeg; ++s) {
        const uint32_t f0 = frameOffsets[s], f1 = frameOffsets[s + 1], n = emOff[s + 1] - emOff[s];
        if (f1 <= f0 || n == 0)
            continue;
        const uint32_t ft0 = ft, nFt = (f1 - f0 + kTileM - 1) / kTileM;
        for (uint32_t k = 0; k < nFt; ++k, ++ft) {  // frame tile ft holds frames [first, first + count)
            tileMap[2 * ft]     = f0 + k * kTileM;
            tileMap[2 * ft + 1] = std::min<uint32_t>(kTileM, f1 - f0 - k * kTileM);
        }
        for (uint32_t v = 0; v * mpt < n; ++v, ++vt) {
            vtBase[vt] = v * mpt;
            for (uint32_t j = 0; j < mpt; ++j) {
                const uint32_t k = v * mpt + j;
                const uint32_t m = emList[emOff[s] + (k < n ? k : 0)];  // spare places repeat the first mixture
                if (m >= h->M) {
                    h->setError("subset scoring: mixture index out of range");
                    return GMM_ERR_ARG;
                }
                vtMix[(size_t)vt * mpt + j] = m;
            }
            uint32_t* it = items + (size_t)vt * 8;
            it[0] = vt, it[1] = ft0, it[2] = ft0 + nFt, it[3] = f0, it[4] = f1, it[5] = it[6] = it[7] = 0;
        }
    }
    if ((rc = h->reserve(SCR_TAB, total, &tab)))
        return rc;
    if ((rc = h->reserve(SCR_VTILES, (size_t)nVt * tm.tileBytes, &vtiles)))
        return rc;
    char* d = (char*)tab;
    if ((rc = gmm_copy_tables(h, d, pin, total)))
        return rc;
    if ((rc = h->pinnedCopied(PIN_SUBSET)))
        return rc;
    gather_tiles_kernel<<<(int)std::min<uint32_t>(nVt, (uint32_t)h->smCount * 8), 256, 0, h->stream>>>(
            tm.bTiles, tm.tileBytes, tm.nAtoms, tm.swz, tm.mixCols, mpt, mixPerPart, parts, (const uint32_t*)(d + oMix),
            (const uint32_t*)(d + oBase), nVt, (uint8_t*)vtiles);
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    TcJob job;
    job.T       = FT * kTileM;
    job.FT      = FT;
    job.d_x     = d_x;
    job.rowMap  = (const uint32_t*)(d + oMap);
    job.bTiles  = (const uint8_t*)vtiles;
    job.nTiles  = nVt;
    job.itemTab = (const uint32_t*)(d + oItem);
    job.nItems  = nVt;
    job.scores  = d_scores;
    job.best    = nullptr;
    job.ld      = ld;
    return tc_run(h, job);
}
