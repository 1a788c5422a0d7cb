// Device-resident consumer of the score block: time-synchronous Viterbi over alignment
// automata, reading the [T x ld] emission scores where the scoring kernels left them, so that only the best
// path (4 bytes per frame) crosses PCIe instead of T x M scores.
//
// Replaces, for the unpruned Viterbi case, the search space of the reference's aligner:
//   Search::Aligner::SearchSpace::addStartupHypothesis / expand / activateOrUpdateStateHypothesisViterbi /
//   getAlignmentFsaViterbi   (/root/reference/src/Search/Aligner.cc:132-158, 190-239, 398-437)
// Semantics kept: the start hypothesis has score 0 before the first frame; every arc consumes one frame and costs
// (score + transitionWeight) + emissionScore in f32, in this order; a state keeps the candidate that arrives last
// among equals (`sh.score >= score`); after the last frame the best FINAL state (score + finalWeight, strict <,
// first among equals) is traced back; a segment whose automaton cannot reach a final state in T frames yields an
// empty alignment (all emissions 0xFFFFFFFF, score FLT_MAX).
// Order of candidates: the reference visits the previous frame's hypotheses in their order of activation and the
// arcs of each in Fsa order, so exact f32 ties (between candidates of one state, between final states) are
// resolved by that order -- and with f32 path scores in the thousands such ties are common, not exotic.  Without
// pruning the order is a property of the automaton alone: the list of frame t is the ordered set of arc targets of
// the list of frame t-1.  align_ranked_kernel carries it as a dense rank per state: a target's arrival key is
// rank(source) * A + arc position in the source, the rank of a new hypothesis is the number of smaller FIRST arrival
// keys (a bitmap of keys + prefix pop-count per frame), and once the ranking repeats (the list is a fixed point of
// the map above) it is frozen.  Among equal candidates of a state the largest arrival key wins (`>=` lets the later
// arrival replace), among equal final hypotheses the smallest rank (strict `<` keeps the first).
// Automata too large for that kernel's shared-memory tables take align_generic_kernel: an order-free parallel pass
// that detects whether any minimum was shared, and, only then, a literal single-thread replay of the reference's
// hypothesis lists (ordered_pass); gmm_align_ordered_count reports how many segments were replayed.
//
// One CTA per segment (segments are independent).  align_ranked_kernel keeps the automaton, the state scores and
// the ranks in shared memory, gathers the emission scores of the next frames with asynchronous 4-byte copies
// (K frames in flight) so that the frame loop never waits for HBM, writes one back-pointer per (frame, state), and
// tracks for every state its ancestor at the last 32-frame boundary, so that the traceback is ceil(T/32) dependent
// steps by one thread followed by independent 32-frame walks by one thread each.
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <vector>

#include "gmm_internal.cuh"

namespace {

constexpr uint32_t kNoArc       = 0xFFFFFFFFu;
constexpr uint32_t kBlockFrames = 32;  // traceback block
constexpr int      kFastArcs    = 4;   // ranked kernel: incoming arcs per state kept in registers once the list is frozen

struct AlignSeg {
    uint32_t frame0, frames;  // rows [frame0, frame0 + frames) of the score block
    uint32_t state0, states;  // states [state0, state0 + states) of the concatenated automata
    uint32_t initial;         // segment-local initial state
    uint32_t keyShift;        // arrival key = rank << keyShift | arc position in its source state
    uint64_t back0;           // offset of this segment's back-pointer block [frames x states]
    uint64_t map0;            // offset of this segment's boundary-ancestor block [ceil(frames / 32) x states]
};

struct AlignParams {
    const AlignSeg* seg;
    uint32_t        nSeg;
    const float*    scores;
    size_t          ld;
    const uint32_t* inOff;  // [nStates + 1] incoming arcs of every (global) state
    const uint32_t* inSrc;  // segment-local source state
    const uint32_t* inSK;   // source state << 16 | position of the arc in its source's arc list (ranked kernel)
    const uint32_t* inEmSlot;  // ranked kernel: position of the arc's emission in its segment's emission list
    const uint32_t* emOff;     // [nSeg + 1] the distinct emission indices each segment's automaton uses ...
    const uint32_t* emList;    // ... (a frame's scores are gathered once per distinct emission, not per arc)
    const uint32_t* inEm;
    const uint32_t* inCol;      // column of the score block an arc reads: inEm, or inEmSlot for a compact block
    uint32_t        compact;    // score block holds, per segment, one column per entry of its emission list
    const float*    inW;
    const uint32_t* outOff;     // [nStates + 1] the caller's arc lists (Fsa order) ...
    const uint32_t* outTarget;  // ... segment-local target ...
    const uint32_t* outToIn;    // ... and the arc's position in the in* arrays
    const float*    finalW;     // [nStates], +inf: not final
    uint32_t*       back;       // back-pointers: index into the in* arrays, or kNoArc
    uint32_t*       map;        // ranked kernel: ancestor state at the last 32-frame boundary, per block and state
    uint32_t*       lists;      // ordered pass: 6 arrays of maxStates words per CTA
    uint32_t        maxStates;
    uint32_t*       orderedCount;
    uint32_t*       pathEmission;
    uint32_t*       pathState;
    float*          pathScore;
    uint32_t        slotOutput;  // pathEmission receives the arc's position in its segment's emission list, not the id
};

// The reference's search space, literally (Aligner.cc:132-158 activateOrUpdateStateHypothesisViterbi, 190-239
// expand, 398-411 best final hypothesis): hypotheses of a frame in activation order, a state's hypothesis is
// replaced when `old >= new`.  One thread.  Returns the best final state (kNoArc: none) and its score.
__device__ uint32_t ordered_pass(const AlignParams& p, const AlignSeg& sg, float* bestScore) {
    const uint32_t S        = sg.states;
    uint32_t*      base     = p.lists + (size_t)blockIdx.x * 6 * p.maxStates;
    uint32_t*      prevSt   = base;
    uint32_t*      curSt    = base + p.maxStates;
    float*         prevSc   = (float*)(base + 2 * (size_t)p.maxStates);
    float*         curSc    = (float*)(base + 3 * (size_t)p.maxStates);
    uint32_t*      stamp    = base + 4 * (size_t)p.maxStates;  // frame + 1 in which pos[] was set
    uint32_t*      pos      = base + 5 * (size_t)p.maxStates;
    uint32_t       nPrev    = 1;
    for (uint32_t s = 0; s < S; ++s)
        stamp[s] = 0;
    prevSt[0] = sg.initial;
    prevSc[0] = 0.f;
    for (uint32_t t = 0; t < sg.frames; ++t) {
        const float* row  = p.scores + (size_t)(sg.frame0 + t) * p.ld;
        uint32_t*    bk   = p.back + sg.back0 + (size_t)t * S;
        uint32_t     nCur = 0;
        for (uint32_t i = 0; i < nPrev; ++i) {
            const uint32_t s  = prevSt[i];
            const float    sc = prevSc[i];
            for (uint32_t a = p.outOff[sg.state0 + s]; a < p.outOff[sg.state0 + s + 1]; ++a) {
                const uint32_t k   = p.outToIn[a];
                const uint32_t tgt = p.outTarget[a];
                const float    c   = __fadd_rn(__fadd_rn(sc, p.inW[k]), row[p.inCol[k]]);
                if (stamp[tgt] != t + 1) {
                    stamp[tgt]  = t + 1;
                    pos[tgt]    = nCur;
                    curSt[nCur] = tgt;
                    curSc[nCur] = c;
                    bk[tgt]     = k;
                    ++nCur;
                }
                else if (curSc[pos[tgt]] >= c) {
                    curSc[pos[tgt]] = c;
                    bk[tgt]         = k;
                }
            }
        }
        uint32_t* ts = prevSt;
        prevSt       = curSt;
        curSt        = ts;
        float* tf    = prevSc;
        prevSc       = curSc;
        curSc        = tf;
        nPrev        = nCur;
    }
    float    best      = FLT_MAX;
    uint32_t bestState = kNoArc;
    for (uint32_t i = 0; i < nPrev; ++i) {
        const float fw = p.finalW[sg.state0 + prevSt[i]];
        if (fw == INFINITY)
            continue;
        const float sc = __fadd_rn(fw, prevSc[i]);
        if (sc < best) {
            best      = sc;
            bestState = prevSt[i];
        }
    }
    *bestScore = best;
    return bestState;
}

__global__ void __launch_bounds__(256) align_generic_kernel(const AlignParams p) {
    extern __shared__ float sm[];
    for (uint32_t si = blockIdx.x; si < p.nSeg; si += gridDim.x) {
        const AlignSeg sg  = p.seg[si];
        const uint32_t S   = sg.states;
        float*         cur = sm;
        float*         nxt = sm + S;
        int            tie = 0;
        for (uint32_t s = threadIdx.x; s < S; s += blockDim.x)
            cur[s] = (s == sg.initial) ? 0.f : INFINITY;
        __syncthreads();
        for (uint32_t t = 0; t < sg.frames; ++t) {
            const float* row = p.scores + (size_t)(sg.frame0 + t) * p.ld;
            uint32_t*    bk  = p.back + sg.back0 + (size_t)t * S;
            for (uint32_t s = threadIdx.x; s < S; s += blockDim.x) {
                const uint32_t a0 = p.inOff[sg.state0 + s], a1 = p.inOff[sg.state0 + s + 1];
                float          best    = INFINITY;
                uint32_t       bestArc = kNoArc;
                int            shared  = 0;  // minimum reached by more than one candidate
                for (uint32_t a = a0; a < a1; ++a) {
                    const float from = cur[p.inSrc[a]];
                    if (from == INFINITY)
                        continue;  // no hypothesis, or one that cannot win any more
                    const float c = __fadd_rn(__fadd_rn(from, p.inW[a]), row[p.inCol[a]]);
                    if (bestArc == kNoArc || c < best) {
                        best    = c;
                        bestArc = a;
                        shared  = 0;
                    }
                    else if (c == best && c < INFINITY)
                        shared = 1;
                }
                tie |= shared;
                nxt[s] = bestArc == kNoArc ? INFINITY : best;
                bk[s]  = bestArc;
            }
            __syncthreads();
            float* sw = cur;
            cur       = nxt;
            nxt       = sw;
        }
        tie = __syncthreads_or(tie);
        if (threadIdx.x == 0) {
            float    best      = FLT_MAX;
            uint32_t bestState = kNoArc;
            for (uint32_t s = 0; s < S && !tie; ++s) {
                const float fw = p.finalW[sg.state0 + s];
                if (cur[s] == INFINITY || fw == INFINITY)
                    continue;
                const float sc = __fadd_rn(fw, cur[s]);
                if (sc < best) {
                    best      = sc;
                    bestState = s;
                }
                else if (sc == best)
                    tie = 1;
            }
            if (tie) {
                bestState = ordered_pass(p, sg, &best);
                atomicAdd(p.orderedCount, 1u);
            }
            p.pathScore[si] = best;
            uint32_t s = bestState;
            for (uint32_t t = sg.frames; t-- > 0;) {
                uint32_t em = kNoArc, st = kNoArc;
                if (s != kNoArc) {
                    const uint32_t a = p.back[sg.back0 + (size_t)t * S + s];
                    st               = s;
                    em               = p.slotOutput ? p.inEmSlot[a] : p.inEm[a];
                    s                = p.inSrc[a];
                }
                p.pathEmission[sg.frame0 + t] = em;
                if (p.pathState)
                    p.pathState[sg.frame0 + t] = st;
            }
        }
        __syncthreads();
    }
}

__device__ __forceinline__ void cp_async4(float* smemDst, const float* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smemDst)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template<int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Shared-memory words align_ranked_kernel needs for an automaton of S states, nIn arcs and W key-bitmap words.
__host__ __device__ inline size_t ranked_smem_words(uint32_t S, uint32_t nIn, uint32_t nEm, uint32_t W, int K) {
    return (size_t)6 * S + (S + 1) + (size_t)3 * nIn + (size_t)(K + 1) * nEm + (size_t)3 * W;
}

template<int K>
__global__ void __launch_bounds__(1024) align_ranked_kernel(const AlignParams p) {
    extern __shared__ uint32_t smw[];
    const uint32_t tid = threadIdx.x, nThreads = blockDim.x;
    for (uint32_t si = blockIdx.x; si < p.nSeg; si += gridDim.x) {
        const AlignSeg sg      = p.seg[si];
        const uint32_t S       = sg.states, T = sg.frames;
        const uint32_t arcBase = p.inOff[sg.state0];
        const uint32_t nIn     = p.inOff[sg.state0 + S] - arcBase;
        const uint32_t W       = (uint32_t)((((uint64_t)S << sg.keyShift) + 31) >> 5);
        float*         cur     = (float*)smw;
        float*         nxt     = cur + S;
        uint32_t*      rankO   = smw + 2 * (size_t)S;
        uint32_t*      rankN   = rankO + S;
        uint32_t*      ancO    = rankN + S;
        uint32_t*      ancN    = ancO + S;
        uint32_t*      offs    = ancN + S;          // [S + 1] incoming arcs of state s: offs[s] .. offs[s + 1]
        uint32_t*      tabSK   = offs + S + 1;      // [nIn]
        float*         tabW    = (float*)(tabSK + nIn);
        uint32_t*      tabEm   = (uint32_t*)(tabW + nIn);  // position in emIds / in a ring row
        const uint32_t em0     = p.emOff[si], nEm = p.emOff[si + 1] - em0;
        uint32_t*      emIds   = tabEm + nIn;               // [nEm] distinct emission indices of this automaton
        float*         ring    = (float*)(emIds + nEm);     // [K][nEm] their scores for the frames in flight
        uint32_t*      bm0     = (uint32_t*)(ring + (size_t)K * nEm);
        uint32_t*      bm1     = bm0 + W;
        uint32_t*      pref    = bm1 + W;
        for (uint32_t s = tid; s <= S; s += nThreads)
            offs[s] = p.inOff[sg.state0 + s] - arcBase;
        for (uint32_t a = tid; a < nIn; a += nThreads) {
            tabSK[a] = p.inSK[arcBase + a];
            tabW[a]  = p.inW[arcBase + a];
            tabEm[a] = p.inEmSlot[arcBase + a];
        }
        for (uint32_t e = tid; e < nEm; e += nThreads)
            emIds[e] = p.emList[em0 + e];
        for (uint32_t s = tid; s < S; s += nThreads) {
            cur[s]   = (s == sg.initial) ? 0.f : INFINITY;
            rankO[s] = (s == sg.initial) ? 0u : kNoArc;
            ancO[s]  = kNoArc;
        }
        for (uint32_t w = tid; w < 2 * W; w += nThreads)
            bm0[w] = 0;
        __syncthreads();
        auto prefetch = [&](uint32_t f) {
            if (f < T) {
                const float* row = p.scores + (size_t)(sg.frame0 + f) * p.ld;
                float*       dst = ring + (size_t)(f % K) * nEm;
                for (uint32_t e = tid; e < nEm; e += nThreads)
                    cp_async4(dst + e, row + (p.compact ? e : emIds[e]));
            }
            cp_async_commit();
        };
        for (uint32_t f = 0; f + 1 < (uint32_t)K; ++f)
            prefetch(f);
        // fast frames (below): one state per thread and no state with more than kFastArcs incoming arcs
        const uint32_t rArc0 = tid < S ? offs[tid] : 0, rDeg = tid < S ? offs[tid + 1] - offs[tid] : 0;
        const int      fast  = __syncthreads_and(rDeg <= (uint32_t)kFastArcs) && S <= nThreads;
        uint32_t       rSrc[kFastArcs], rEm[kFastArcs], rKey[kFastArcs];
        float          rW[kFastArcs];
#pragma unroll
        for (int j = 0; j < kFastArcs; ++j) {
            rSrc[j] = rEm[j] = 0;
            rKey[j] = kNoArc;
            rW[j]   = 0.f;
        }
        int stable = 0;
        for (uint32_t t = 0; t < T; ++t) {
            cp_async_wait<K - 2>();
            __syncthreads();  // frame t's scores landed; frame t-1's state scores, ranks and ancestors are complete
            prefetch(t + K - 1);
            const float*   E        = ring + (size_t)(t % K) * nEm;
            uint32_t*      bm       = (t & 1) ? bm1 : bm0;
            uint32_t*      bk       = p.back + sg.back0 + (size_t)t * S;
            const bool     boundary = (t % kBlockFrames) == 0;
            const bool     emitMap  = ((t + 1) % kBlockFrames) == 0 || t + 1 == T;
            uint32_t*      mapRow   = p.map + sg.map0 + (size_t)(t / kBlockFrames) * S;
            const uint32_t keyMask  = (1u << sg.keyShift) - 1;
            if (stable && fast) {
                // frozen hypothesis list, one state per thread, its (at most 4) incoming arcs in registers: a frame is
                // two rounds of independent shared-memory loads, the comparisons, three stores and one barrier
                if (tid < S) {
                    float    best = INFINITY;
                    uint32_t bestJ = kNoArc, bestKey = 0, bestSrc = 0;
#pragma unroll
                    for (int j = 0; j < kFastArcs; ++j) {
                        if (rKey[j] != kNoArc) {  // the arc exists and its source carries a hypothesis
                            const float c = __fadd_rn(__fadd_rn(cur[rSrc[j]], rW[j]), E[rEm[j]]);
                            if (bestJ == kNoArc || c < best || (c == best && rKey[j] > bestKey)) {
                                best    = c;
                                bestJ   = (uint32_t)j;
                                bestKey = rKey[j];
                                bestSrc = rSrc[j];
                            }
                        }
                    }
                    nxt[tid] = best;
                    bk[tid]  = bestJ == kNoArc ? kNoArc : rArc0 + bestJ;
                    const uint32_t anc = bestJ == kNoArc ? kNoArc : (boundary ? bestSrc : ancO[bestSrc]);
                    ancN[tid]          = anc;
                    if (emitMap)
                        mapRow[tid] = anc;
                }
            }
            else
            for (uint32_t s = tid; s < S; s += nThreads) {
                float    best = INFINITY;
                uint32_t bestArc = kNoArc, bestKey = 0, firstKey = kNoArc, bestSrc = 0;
                const uint32_t a1 = offs[s + 1];
#pragma unroll 4
                for (uint32_t a = offs[s]; a < a1; ++a) {  // no early exit: the loads of several arcs overlap
                    const uint32_t sk = tabSK[a], src = sk >> 16;
                    const uint32_t r   = rankO[src];  // kNoArc: no hypothesis in the previous frame
                    const float    c   = __fadd_rn(__fadd_rn(cur[src], tabW[a]), E[tabEm[a]]);
                    const uint32_t key = (r << sg.keyShift) | (sk & keyMask);
                    const bool     hyp = r != kNoArc;
                    firstKey           = hyp ? min(firstKey, key) : firstKey;
                    if (hyp && (bestArc == kNoArc || c < best || (c == best && key > bestKey))) {
                        best    = c;
                        bestArc = a;
                        bestKey = key;
                        bestSrc = src;
                    }
                }
                nxt[s] = best;
                bk[s]  = bestArc;
                const uint32_t anc = bestArc == kNoArc ? kNoArc : (boundary ? bestSrc : ancO[bestSrc]);
                ancN[s]            = anc;
                if (emitMap)
                    mapRow[s] = anc;
                if (!stable) {
                    rankN[s] = firstKey;
                    if (firstKey != kNoArc)
                        atomicOr(&bm[firstKey >> 5], 1u << (firstKey & 31));
                }
            }
            if (!stable) {
                __syncthreads();
                if (tid < 32) {  // exclusive prefix pop-count of the key bitmap
                    uint32_t carry = 0;
                    for (uint32_t w0 = 0; w0 < W; w0 += 32) {
                        const uint32_t w = w0 + tid;
                        const uint32_t v = w < W ? __popc(bm[w]) : 0;
                        uint32_t       x = v;
                        for (int d = 1; d < 32; d <<= 1) {
                            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, d);
                            if ((int)tid >= d)
                                x += y;
                        }
                        if (w < W)
                            pref[w] = carry + x - v;
                        carry += __shfl_sync(0xFFFFFFFFu, x, 31);
                    }
                }
                __syncthreads();
                int same = 1;
                for (uint32_t s = tid; s < S; s += nThreads) {
                    const uint32_t k = rankN[s];
                    uint32_t       r = kNoArc;
                    if (k != kNoArc)
                        r = pref[k >> 5] + __popc(bm[k >> 5] & ((1u << (k & 31)) - 1));
                    rankN[s] = r;
                    same &= (r == rankO[s]);
                }
                stable = __syncthreads_and(same);  // the hypothesis list repeats: it is a fixed point from here on
                for (uint32_t w = tid; w < W; w += nThreads)
                    bm[w] = 0;
                uint32_t* sw = rankO;
                rankO        = rankN;
                rankN        = sw;
                if (stable && fast && tid < S) {  // the frozen ranks give every arc its final arrival key
#pragma unroll
                    for (int j = 0; j < kFastArcs; ++j) {
                        rKey[j] = kNoArc;
                        if ((uint32_t)j < rDeg) {
                            const uint32_t sk = tabSK[rArc0 + j], r = rankO[sk >> 16];
                            rSrc[j]           = sk >> 16;
                            rW[j]             = tabW[rArc0 + j];
                            rEm[j]            = tabEm[rArc0 + j];
                            rKey[j]           = r == kNoArc ? kNoArc : ((r << sg.keyShift) | (sk & keyMask));
                        }
                    }
                }
            }
            float* sf = cur;
            cur       = nxt;
            nxt       = sf;
            uint32_t* sa = ancO;
            ancO         = ancN;
            ancN         = sa;
        }
        cp_async_wait<0>();
        __syncthreads();
        // best final hypothesis: smallest score, the earliest activated among equals (Aligner.cc:398-411)
        const uint32_t nBlocks = (T + kBlockFrames - 1) / kBlockFrames;
        uint32_t*      dState  = p.pathState + sg.frame0;
        __shared__ uint32_t bestShared;
        __shared__ float    redScore[32];
        __shared__ uint32_t redRank[32], redState[32];
        {
            float    best = FLT_MAX;
            uint32_t bestState = kNoArc, bestRank = kNoArc;
            for (uint32_t s = tid; s < S; s += nThreads) {
                const float fw = p.finalW[sg.state0 + s];
                if (rankO[s] == kNoArc || fw == INFINITY)
                    continue;
                const float sc = __fadd_rn(fw, cur[s]);
                if (sc < best || (sc == best && sc < FLT_MAX && rankO[s] < bestRank)) {
                    best      = sc;
                    bestState = s;
                    bestRank  = rankO[s];
                }
            }
            auto merge = [&](float sc, uint32_t rk, uint32_t st) {
                if (sc < best || (sc == best && rk < bestRank)) {
                    best      = sc;
                    bestRank  = rk;
                    bestState = st;
                }
            };
            for (int d = 16; d > 0; d >>= 1)
                merge(__shfl_xor_sync(0xFFFFFFFFu, best, d), __shfl_xor_sync(0xFFFFFFFFu, bestRank, d),
                      __shfl_xor_sync(0xFFFFFFFFu, bestState, d));
            if ((tid & 31) == 0) {
                redScore[tid >> 5] = best;
                redRank[tid >> 5]  = bestRank;
                redState[tid >> 5] = bestState;
            }
        }
        __syncthreads();
        if (tid == 0) {
            float    best = FLT_MAX;
            uint32_t bestState = kNoArc, bestRank = kNoArc;
            for (uint32_t w = 0; w < (nThreads + 31) / 32; ++w)
                if (redScore[w] < best || (redScore[w] == best && redRank[w] < bestRank)) {
                    best      = redScore[w];
                    bestRank  = redRank[w];
                    bestState = redState[w];
                }
            p.pathScore[si] = best;
            bestShared      = bestState;
            // state after the last frame of every 32-frame block, walking the boundary-ancestor maps backwards
            uint32_t s = bestState;
            for (uint32_t b = nBlocks; b-- > 0 && s != kNoArc;) {
                const uint32_t tLast = min(T, (b + 1) * kBlockFrames) - 1;
                dState[tLast]        = s;
                s                    = p.map[sg.map0 + (size_t)b * S + s];
            }
        }
        __syncthreads();
        if (bestShared == kNoArc) {
            for (uint32_t t = tid; t < T; t += nThreads) {
                p.pathEmission[sg.frame0 + t] = kNoArc;
                dState[t]                     = kNoArc;
            }
        }
        else {
            for (uint32_t b = tid; b < nBlocks; b += nThreads) {
                const uint32_t tFirst = b * kBlockFrames, tLast = min(T, tFirst + kBlockFrames) - 1;
                uint32_t       s      = dState[tLast];
                for (uint32_t t = tLast + 1; t-- > tFirst;) {
                    const uint32_t a              = p.back[sg.back0 + (size_t)t * S + s];
                    dState[t]                     = s;
                    p.pathEmission[sg.frame0 + t] = p.slotOutput ? (uint32_t)tabEm[a] : emIds[tabEm[a]];
                    s                             = tabSK[a] >> 16;
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace

bool gmm_align_emission_lists(gmm_handle* h, uint32_t nSeg, const uint32_t* stateOffsets, const uint32_t* arcOffsets,
                              const uint32_t* arcEmission, uint32_t nEmissions, AlignLists& out) {
    for (uint32_t i = 0; i < nSeg; ++i)
        if (stateOffsets[i + 1] < stateOffsets[i] || arcOffsets[stateOffsets[i + 1]] < arcOffsets[stateOffsets[i]])
            return false;
    const uint32_t nArcs = arcOffsets[stateOffsets[nSeg]] - arcOffsets[stateOffsets[0]], arc0 = arcOffsets[stateOffsets[0]];
    out.arcSlot.resize(nArcs);
    out.emOff.assign(nSeg + 1, 0);
    std::vector<uint32_t>& sparse = out.scratch;  // segment i's list at its first arc; emOff[i + 1] = its length
    sparse.resize(nArcs);
    std::atomic<bool> bad(false);
    HostPool&         pool = h->hostPool();
    pool.run(nSeg, std::max(1u, nSeg / (4 * (uint32_t)pool.size())), [&](uint32_t i0, uint32_t i1, int worker) {
        gmm_handle::EmissionStamps& w = h->workerStamps[worker];
        if (w.stamp.size() != nEmissions) {
            w.slotOf.assign(nEmissions, 0);
            w.stamp.assign(nEmissions, 0);
            w.counter = 0;
        }
        uint32_t *slotOf = w.slotOf.data(), *stamp = w.stamp.data(), &counter = w.counter;
        for (uint32_t i = i0; i < i1; ++i) {
            if (++counter == 0) {  // stamps wrapped: start over
                std::fill(stamp, stamp + nEmissions, 0u);
                counter = 1;
            }
            const uint32_t a0 = arcOffsets[stateOffsets[i]], a1 = arcOffsets[stateOffsets[i + 1]];
            uint32_t*      list = sparse.data() + (a0 - arc0);
            uint32_t       n    = 0;
            for (uint32_t a = a0; a < a1; ++a) {
                const uint32_t e = arcEmission[a];
                if (e >= nEmissions) {
                    bad = true;
                    break;
                }
                if (stamp[e] != counter) {  // first use of this emission in segment i
                    stamp[e]  = counter;
                    slotOf[e] = n;
                    list[n++] = e;
                }
                out.arcSlot[a - arc0] = slotOf[e];
            }
            out.emOff[i + 1] = n;
        }
    });
    if (bad)
        return false;
    for (uint32_t i = 0; i < nSeg; ++i)
        out.emOff[i + 1] += out.emOff[i];
    out.emList.resize(out.emOff[nSeg]);
    for (uint32_t i = 0; i < nSeg; ++i)
        if (out.emOff[i + 1] > out.emOff[i])
            memcpy(out.emList.data() + out.emOff[i], sparse.data() + (arcOffsets[stateOffsets[i]] - arc0),
                   (size_t)(out.emOff[i + 1] - out.emOff[i]) * 4);
    return true;
}

// Host arrays in, host arrays out; `d_scores` is the DEVICE score block.  See include/gmm_b200.h.
int gmm_align_launch(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                     const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                     const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                     const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState, float* pathScore,
                     const uint32_t** d_pathEmission, bool compact, uint32_t nEmissions, AlignDeferred* defer,
                     const AlignLists* lists, bool slotOutput) {
    if (defer)
        defer->pending = false;
    if (d_pathEmission)
        *d_pathEmission = nullptr;
    if (nSeg == 0)
        return GMM_OK;
    const uint32_t nStates = stateOffsets[nSeg], nFrames = frameOffsets[nSeg];
    const uint32_t nArcs   = arcOffsets[nStates];
    // the distinct emissions of every automaton (one gather / one compact column each) and each arc's place in them
    AlignLists        own;
    const AlignLists* L = lists;
    const uint32_t    emLimit = compact ? nEmissions : (uint32_t)std::min<size_t>(ld, 0xFFFFFFFFu);
    if (!L) {
        if (!gmm_align_emission_lists(h, nSeg, stateOffsets, arcOffsets, arcEmission, emLimit, own)) {
            h->setError("gmm_align: arc offsets not ascending or emission index outside the score block");
            return GMM_ERR_ARG;
        }
        L = &own;
    }
    const std::vector<uint32_t>&emOff = L->emOff, &emList = L->emList, &arcSlot = L->arcSlot;
    // Every table the kernels read is built in ONE pinned host block laid out like its device copy (tables of both
    // kernels first, those only the generic kernel reads last), so a launch costs one asynchronous copy.
    auto   al   = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t off  = 0;
    auto   take = [&](size_t bytes) {
        size_t o = off;
        off += al(bytes);
        return o;
    };
    const size_t oSeg = take(nSeg * sizeof(AlignSeg)), oOff = take((nStates + 1) * 4), oSK = take((size_t)nArcs * 4),
                 oEm = take((size_t)nArcs * 4), oW = take((size_t)nArcs * 4), oFin = take((size_t)nStates * 4),
                 oSlot = take((size_t)nArcs * 4), oEmOff = take((nSeg + 1) * 4), oEmList = take(emList.size() * 4),
                 oCnt = take(4), rankedBytes = off, oSrc = take((size_t)nArcs * 4), oOut = take((nStates + 1) * 4),
                 oTgt = take((size_t)nArcs * 4), oO2i = take((size_t)nArcs * 4), tableBytes = off;
    void* pin = nullptr;
    int   rc  = h->reservePinned(PIN_ALIGN, tableBytes, &pin);
    if (rc)
        return rc;
    char*     hp       = (char*)pin;
    AlignSeg* segs     = (AlignSeg*)(hp + oSeg);
    uint32_t *inOff = (uint32_t*)(hp + oOff), *inSK = (uint32_t*)(hp + oSK), *inEm = (uint32_t*)(hp + oEm),
             *inEmSlot = (uint32_t*)(hp + oSlot), *inSrc = (uint32_t*)(hp + oSrc), *outToIn = (uint32_t*)(hp + oO2i);
    float*    inW      = (float*)(hp + oW);
    *(uint32_t*)(hp + oCnt) = 0;
    // segments are independent (targets are segment-local: the incoming arcs of a segment's states are its own arcs,
    // so its part of the incoming-arc CSR starts where its arcs start); the per-arc work runs on the host pool
    uint64_t backTotal = 0, mapTotal = 0;
    uint32_t maxStates = 0;
    for (uint32_t i = 0; i < nSeg; ++i) {
        const uint32_t s0 = stateOffsets[i], s1 = stateOffsets[i + 1];
        if (s1 < s0 || frameOffsets[i + 1] < frameOffsets[i] || initialState[i] >= std::max(1u, s1 - s0)) {
            h->setError("gmm_align: offsets not ascending or initial state out of range");
            return GMM_ERR_ARG;
        }
        if (compact && emOff[i + 1] - emOff[i] > ld) {
            h->setError("gmm_align: compact score block narrower than a segment's emission list");
            return GMM_ERR_ARG;
        }
        AlignSeg& sg = segs[i];
        sg.frame0    = frameOffsets[i];
        sg.frames    = frameOffsets[i + 1] - frameOffsets[i];
        sg.state0    = s0;
        sg.states    = s1 - s0;
        sg.initial   = initialState[i];
        sg.keyShift  = 0;
        sg.back0     = backTotal;
        sg.map0      = mapTotal;
        backTotal += (uint64_t)sg.frames * sg.states;
        mapTotal += (uint64_t)((sg.frames + kBlockFrames - 1) / kBlockFrames) * sg.states;
        maxStates = std::max(maxStates, sg.states);
    }
    const bool forceGeneric = [] {
        const char* f = getenv("GMM_ALIGN_KERNEL");
        return f && !strcmp(f, "generic");
    }();
    HostPool&             pool = h->hostPool();
    std::vector<uint32_t> segDegree(nSeg, 0);
    std::atomic<int>      bad(0);
    inOff[0] = 0;
    pool.run(nSeg, std::max(1u, nSeg / (4 * (uint32_t)pool.size())), [&](uint32_t i0, uint32_t i1, int worker) {
        std::vector<uint32_t>& fill = h->workerWords[worker];  // next free place of every state's list
        if (fill.size() < (size_t)maxStates + 1)
            fill.resize((size_t)maxStates + 1);
        for (uint32_t i = i0; i < i1; ++i) {
            const uint32_t st0 = stateOffsets[i], S = stateOffsets[i + 1] - st0, arcBase = arcOffsets[st0];
            uint32_t       degree = 0;
            std::fill(fill.begin(), fill.begin() + S + 1, 0u);
            for (uint32_t s = st0; s < st0 + S; ++s) {
                if (arcOffsets[s + 1] < arcOffsets[s]) {
                    bad = 1;
                    return;
                }
                degree = std::max(degree, arcOffsets[s + 1] - arcOffsets[s]);
                for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a) {
                    if (arcTarget[a] >= S) {
                        bad = 2;
                        return;
                    }
                    fill[arcTarget[a] + 1]++;
                }
            }
            uint32_t run = arcBase;  // counts -> first place of every state's list
            for (uint32_t s = 0; s < S; ++s) {
                const uint32_t c   = fill[s + 1];
                fill[s]            = run;
                run += c;
                inOff[st0 + s + 1] = run;
            }
            for (uint32_t s = st0; s < st0 + S; ++s)  // ascending source, then arc index
                for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a) {
                    const uint32_t k = fill[arcTarget[a]]++;
                    inSK[k]          = ((s - st0) << 16) | ((a - arcOffsets[s]) & 0xFFFFu);
                    inEm[k]          = arcEmission[a];
                    inEmSlot[k]      = arcSlot[a];
                    inW[k]           = arcWeight[a];
                    if (forceGeneric) {
                        inSrc[k]   = s - st0;
                        outToIn[a] = k;
                    }
                }
            segDegree[i] = degree;
        }
    });
    if (bad) {
        h->setError(bad == 1 ? "gmm_align: arc offsets not ascending" : "gmm_align: arc target outside its segment's automaton");
        return GMM_ERR_ARG;
    }
    uint32_t maxDegree = 0;
    size_t   maxWords[3] = {0, 0, 0};  // shared memory of the ranked kernel with 8 / 4 / 2 frames in flight
    for (uint32_t i = 0; i < nSeg; ++i) {
        uint32_t shift = 0;
        while ((1u << shift) < segDegree[i])
            ++shift;
        segs[i].keyShift = shift;
        maxDegree        = std::max(maxDegree, segDegree[i]);
        const uint64_t keyBits = (uint64_t)segs[i].states << shift;
        const uint32_t segArcs = arcOffsets[stateOffsets[i + 1]] - arcOffsets[stateOffsets[i]];
        for (int k = 0; k < 3; ++k)
            maxWords[k] = std::max(maxWords[k], keyBits >= ((uint64_t)1 << 31)
                                                        ? (size_t)1 << 40
                                                        : ranked_smem_words(segs[i].states, segArcs, emOff[i + 1] - emOff[i],
                                                                            (uint32_t)((keyBits + 31) >> 5), 8 >> k));
    }
    // kernel choice: the ranked kernel when every automaton's tables fit in shared memory
    int inFlight = 0;  // 0: generic kernel
    if (!forceGeneric && maxStates < 65536 && maxDegree < 65536) {
        if (maxWords[0] * 4 <= 110 * 1024)
            inFlight = 8;
        else if (maxWords[1] * 4 <= 200 * 1024)
            inFlight = 4;
        else if (maxWords[2] * 4 <= 200 * 1024)
            inFlight = 2;
    }
    const size_t smemBytes = inFlight ? maxWords[inFlight == 8 ? 0 : inFlight == 4 ? 1 : 2] * 4 + 16
                                      : (size_t)2 * maxStates * sizeof(float);
    if (smemBytes > 200 * 1024) {
        h->setError("gmm_align: automaton with more than 25600 states");
        return GMM_ERR_UNSUPPORTED;
    }
    if (!inFlight && !forceGeneric)  // the generic kernel's own tables (rare: automata beyond the ranked kernel)
        for (uint32_t k = 0; k < nArcs; ++k)
            inSrc[k] = inSK[k] >> 16;
    if (!inFlight && !forceGeneric) {
        std::vector<uint32_t> fill(inOff, inOff + nStates);
        for (uint32_t i = 0; i < nSeg; ++i)
            for (uint32_t s = stateOffsets[i]; s < stateOffsets[i + 1]; ++s)
                for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a)
                    outToIn[a] = fill[stateOffsets[i] + arcTarget[a]]++;
    }
    memcpy(hp + oFin, finalWeight, (size_t)nStates * 4);
    memcpy(hp + oEmOff, emOff.data(), (size_t)(nSeg + 1) * 4);
    if (!emList.empty())
        memcpy(hp + oEmList, emList.data(), emList.size() * 4);
    if (!inFlight) {
        memcpy(hp + oOut, arcOffsets, (size_t)(nStates + 1) * 4);
        if (nArcs)
            memcpy(hp + oTgt, arcTarget, (size_t)nArcs * 4);
    }
    const int grid = (int)std::min<uint32_t>(nSeg, (uint32_t)h->smCount * 8);
    // device block: the tables, then outputs, back-pointers and per-kernel work space
    const size_t oPe = take((size_t)nFrames * 4), oPs = take((size_t)nFrames * 4), oSc = take((size_t)nSeg * 4),
                 oBack = take(backTotal * 4), oMap = take(inFlight ? mapTotal * 4 : 0),
                 oLists = take(inFlight ? 0 : (size_t)grid * 6 * maxStates * 4);
    void* base = nullptr;
    if ((rc = h->reserve(SCR_MISC, off, &base)))
        return rc;
    char* d = (char*)base;
    if ((rc = gmm_copy_tables(h, d, hp, inFlight ? rankedBytes : tableBytes)))
        return rc;
    if ((rc = h->pinnedCopied(PIN_ALIGN)))
        return rc;
    AlignParams p;
    p.seg          = (const AlignSeg*)(d + oSeg);
    p.nSeg         = nSeg;
    p.scores       = d_scores;
    p.ld           = ld;
    p.inOff        = (const uint32_t*)(d + oOff);
    p.inSrc        = (const uint32_t*)(d + oSrc);
    p.inSK         = (const uint32_t*)(d + oSK);
    p.inEm         = (const uint32_t*)(d + oEm);
    p.inEmSlot     = (const uint32_t*)(d + oSlot);
    p.inCol        = compact ? p.inEmSlot : p.inEm;
    p.compact      = compact ? 1u : 0u;
    p.emOff        = (const uint32_t*)(d + oEmOff);
    p.emList       = (const uint32_t*)(d + oEmList);
    p.inW          = (const float*)(d + oW);
    p.outOff       = (const uint32_t*)(d + oOut);
    p.outTarget    = (const uint32_t*)(d + oTgt);
    p.outToIn      = (const uint32_t*)(d + oO2i);
    p.finalW       = (const float*)(d + oFin);
    p.back         = (uint32_t*)(d + oBack);
    p.map          = (uint32_t*)(d + oMap);
    p.lists        = (uint32_t*)(d + oLists);
    p.maxStates    = maxStates;
    p.orderedCount = (uint32_t*)(d + oCnt);
    p.pathEmission = (uint32_t*)(d + oPe);
    p.pathState    = (uint32_t*)(d + oPs);
    p.pathScore    = (float*)(d + oSc);
    p.slotOutput   = slotOutput ? 1u : 0u;
    void (*kernel)(AlignParams) = inFlight == 8   ? align_ranked_kernel<8>
                                  : inFlight == 4 ? align_ranked_kernel<4>
                                  : inFlight == 2 ? align_ranked_kernel<2>
                                                  : align_generic_kernel;
    if (smemBytes > 48 * 1024)
        GMM_CUDA_TRY(h, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
    h->timeBegin();
    // ranked kernel: one thread per state where that is possible (the frame loop is a latency chain per thread)
    const int threads = inFlight ? (int)std::min<uint32_t>(1024, std::max<uint32_t>(128, (maxStates + 31) / 32 * 32)) : 256;
    kernel<<<grid, threads, smemBytes, h->stream>>>(p);
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    if (d_pathEmission)
        *d_pathEmission = p.pathEmission;  // stays valid until the next alignment call on this handle
    AlignDeferred now;
    AlignDeferred& fin = defer ? *defer : now;
    fin.pending        = true;
    fin.dPathEmission  = (const uint8_t*)(d + oPe);
    fin.dPathState     = (const uint8_t*)(d + oPs);
    fin.dPathScore     = (const uint8_t*)(d + oSc);
    fin.dOrdered       = (const uint8_t*)(d + oCnt);
    fin.nFrames        = nFrames;
    fin.nSeg           = nSeg;
    fin.pathEmission   = pathEmission;
    fin.pathState      = pathState;
    fin.pathScore      = pathScore;
    return defer ? GMM_OK : gmm_align_finish(h, fin);
}

namespace {
__global__ void __launch_bounds__(256) scale_block_kernel(float* d, size_t rows, size_t cols, size_t ld, float scale) {
    const size_t total = rows * cols;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / cols, c = i - r * cols;
        d[r * ld + c]  = __fmul_rn(scale, d[r * ld + c]);
    }
}
}  // namespace

// Mm::FeatureScorerScaling (ScaledFeatureScorer.hh:62-64): the aligner sees scale_ * score, one f32 product per score.
int gmm_scale_block(gmm_handle* h, float* d, size_t rows, size_t cols, size_t ld, float scale) {
    if (!rows || !cols)
        return GMM_OK;
    const size_t total = rows * cols;
    const int    grid  = (int)std::min<size_t>((total + 255) / 256, (size_t)h->smCount * 8);
    h->timeBegin();
    scale_block_kernel<<<grid, 256, 0, h->stream>>>(d, rows, cols, ld, scale);
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    return GMM_OK;
}

int gmm_align_finish(gmm_handle* h, AlignDeferred& f) {
    if (!f.pending)
        return GMM_OK;
    f.pending = false;
    if (f.pathEmission)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(f.pathEmission, f.dPathEmission, (size_t)f.nFrames * 4, cudaMemcpyDeviceToHost, h->stream));
    if (f.pathState)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(f.pathState, f.dPathState, (size_t)f.nFrames * 4, cudaMemcpyDeviceToHost, h->stream));
    if (f.pathScore)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(f.pathScore, f.dPathScore, (size_t)f.nSeg * 4, cudaMemcpyDeviceToHost, h->stream));
    if (f.weights && f.wRows && f.wCols)
        GMM_CUDA_TRY(h, cudaMemcpy2DAsync(f.weights, f.ldWhost * 4, f.dWeights, f.ldWdev * 4, (size_t)f.wCols * 4, f.wRows,
                                          cudaMemcpyDeviceToHost, h->stream));
    if (f.total)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(f.total, f.dTotal, (size_t)f.nSeg * 8, cudaMemcpyDeviceToHost, h->stream));
    uint32_t ordered = 0;
    if (f.dOrdered)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(&ordered, f.dOrdered, 4, cudaMemcpyDeviceToHost, h->stream));
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    h->alignOrdered += ordered;
    return GMM_OK;
}
