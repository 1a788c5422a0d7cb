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
// resolved by that order.  The parallel pass below is order-free: it finds each state's minimum with one thread per
// state over the state's incoming arcs and records whether any minimum was reached by two candidates.  A segment
// with no such tie has a unique best path whatever the order.  A segment WITH a tie is redone by one thread of the
// same CTA that replays the reference's hypothesis lists literally (ordered_pass), so the result is the
// reference's in both cases; the number of segments that took the ordered pass is reported (gmm_align_ordered_count).
//
// One CTA per segment (segments are independent): state scores of the previous / current frame in shared memory,
// one thread per state and frame, back-pointers [T x S] in HBM, traceback by one thread.
#include <float.h>
#include <math.h>

#include <algorithm>
#include <vector>

#include "gmm_internal.cuh"

namespace {

constexpr uint32_t kNoArc = 0xFFFFFFFFu;

struct AlignSeg {
    uint32_t frame0, frames;  // rows [frame0, frame0 + frames) of the score block
    uint32_t state0, states;  // states [state0, state0 + states) of the concatenated automata
    uint32_t initial;         // segment-local initial state
    uint32_t pad;
    uint64_t back0;           // offset of this segment's back-pointer block [frames x states]
};

struct AlignParams {
    const AlignSeg* seg;
    uint32_t        nSeg;
    const float*    scores;
    size_t          ld;
    const uint32_t* inOff;  // [nStates + 1] incoming arcs of every (global) state
    const uint32_t* inSrc;  // segment-local source state
    const uint32_t* inEm;
    const float*    inW;
    const uint32_t* outOff;     // [nStates + 1] the caller's arc lists (Fsa order) ...
    const uint32_t* outTarget;  // ... segment-local target ...
    const uint32_t* outToIn;    // ... and the arc's position in the in* arrays
    const float*    finalW;     // [nStates], +inf: not final
    uint32_t*       back;       // back-pointers: index into the in* arrays, or kNoArc
    uint32_t*       lists;      // ordered pass: 6 arrays of maxStates words per CTA
    uint32_t        maxStates;
    uint32_t*       orderedCount;
    uint32_t*       pathEmission;
    uint32_t*       pathState;
    float*          pathScore;
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
                const float    c   = __fadd_rn(__fadd_rn(sc, p.inW[k]), row[p.inEm[k]]);
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

__global__ void __launch_bounds__(256) align_kernel(const AlignParams p) {
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
                    const float c = __fadd_rn(__fadd_rn(from, p.inW[a]), row[p.inEm[a]]);
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
                    em               = p.inEm[a];
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

}  // namespace

// Host arrays in, host arrays out; `d_scores` is the DEVICE score block.  See include/gmm_b200.h.
int gmm_align_launch(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                     const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                     const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                     const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState, float* pathScore) {
    if (nSeg == 0)
        return GMM_OK;
    const uint32_t nStates = stateOffsets[nSeg], nFrames = frameOffsets[nSeg];
    const uint32_t nArcs   = arcOffsets[nStates];
    // incoming-arc lists (CSR by target), arcs of a state ordered by (source state, arc index)
    std::vector<uint32_t> inOff(nStates + 1, 0), inSrc(nArcs), inEm(nArcs), outToIn(nArcs);
    std::vector<float>    inW(nArcs);
    std::vector<AlignSeg> segs(nSeg);
    uint64_t              backTotal = 0;
    uint32_t              maxStates = 0;
    for (uint32_t i = 0; i < nSeg; ++i) {
        const uint32_t s0 = stateOffsets[i], s1 = stateOffsets[i + 1];
        if (s1 < s0 || frameOffsets[i + 1] < frameOffsets[i] || initialState[i] >= std::max(1u, s1 - s0)) {
            h->setError("gmm_align: offsets not ascending or initial state out of range");
            return GMM_ERR_ARG;
        }
        for (uint32_t s = s0; s < s1; ++s) {
            if (arcOffsets[s + 1] < arcOffsets[s]) {
                h->setError("gmm_align: arc offsets not ascending");
                return GMM_ERR_ARG;
            }
            for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a) {
                if (arcTarget[a] >= s1 - s0) {
                    h->setError("gmm_align: arc target outside its segment's automaton");
                    return GMM_ERR_ARG;
                }
                if (arcEmission[a] >= ld) {
                    h->setError("gmm_align: emission index >= ld");
                    return GMM_ERR_ARG;
                }
                inOff[s0 + arcTarget[a] + 1]++;
            }
        }
        segs[i].frame0  = frameOffsets[i];
        segs[i].frames  = frameOffsets[i + 1] - frameOffsets[i];
        segs[i].state0  = s0;
        segs[i].states  = s1 - s0;
        segs[i].initial = initialState[i];
        segs[i].pad     = 0;
        segs[i].back0   = backTotal;
        backTotal += (uint64_t)segs[i].frames * segs[i].states;
        maxStates = std::max(maxStates, s1 - s0);
    }
    for (uint32_t s = 0; s < nStates; ++s)
        inOff[s + 1] += inOff[s];
    {
        std::vector<uint32_t> fill(inOff.begin(), inOff.end() - 1);
        for (uint32_t i = 0; i < nSeg; ++i)
            for (uint32_t s = stateOffsets[i]; s < stateOffsets[i + 1]; ++s)  // ascending source, then arc index
                for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a) {
                    const uint32_t k = fill[stateOffsets[i] + arcTarget[a]]++;
                    inSrc[k]         = s - stateOffsets[i];
                    inEm[k]          = arcEmission[a];
                    inW[k]           = arcWeight[a];
                    outToIn[a]       = k;
                }
    }
    const size_t smemBytes = (size_t)2 * maxStates * sizeof(float);
    if (smemBytes > 200 * 1024) {
        h->setError("gmm_align: automaton with more than 25600 states");
        return GMM_ERR_UNSUPPORTED;
    }
    // device buffers: one allocation, carved
    auto   al   = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t off  = 0;
    auto   take = [&](size_t bytes) {
        size_t o = off;
        off += al(bytes);
        return o;
    };
    const size_t oSeg = take(nSeg * sizeof(AlignSeg)), oOff = take((nStates + 1) * 4), oSrc = take((size_t)nArcs * 4),
                 oEm = take((size_t)nArcs * 4), oW = take((size_t)nArcs * 4), oFin = take((size_t)nStates * 4),
                 oPe = take((size_t)nFrames * 4), oPs = take((size_t)nFrames * 4), oSc = take((size_t)nSeg * 4),
                 oOut = take((nStates + 1) * 4), oTgt = take((size_t)nArcs * 4), oO2i = take((size_t)nArcs * 4),
                 oCnt = take(4), oBack = take(backTotal * 4);
    const int    grid   = (int)std::min<uint32_t>(nSeg, (uint32_t)h->smCount * 8);
    const size_t oLists = take((size_t)grid * 6 * maxStates * 4);
    void* base = nullptr;
    int   rc   = h->reserve(SCR_MISC, off, &base);
    if (rc)
        return rc;
    char* d = (char*)base;
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oSeg, segs.data(), nSeg * sizeof(AlignSeg), cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oOff, inOff.data(), (nStates + 1) * 4, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oSrc, inSrc.data(), (size_t)nArcs * 4, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oEm, inEm.data(), (size_t)nArcs * 4, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oW, inW.data(), (size_t)nArcs * 4, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oOut, arcOffsets, (nStates + 1) * 4, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oTgt, arcTarget, (size_t)nArcs * 4, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oO2i, outToIn.data(), (size_t)nArcs * 4, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemsetAsync(d + oCnt, 0, 4, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d + oFin, finalWeight, (size_t)nStates * 4, cudaMemcpyHostToDevice, h->stream));
    AlignParams p;
    p.seg          = (const AlignSeg*)(d + oSeg);
    p.nSeg         = nSeg;
    p.scores       = d_scores;
    p.ld           = ld;
    p.inOff        = (const uint32_t*)(d + oOff);
    p.inSrc        = (const uint32_t*)(d + oSrc);
    p.inEm         = (const uint32_t*)(d + oEm);
    p.inW          = (const float*)(d + oW);
    p.outOff       = (const uint32_t*)(d + oOut);
    p.outTarget    = (const uint32_t*)(d + oTgt);
    p.outToIn      = (const uint32_t*)(d + oO2i);
    p.finalW       = (const float*)(d + oFin);
    p.lists        = (uint32_t*)(d + oLists);
    p.maxStates    = maxStates;
    p.orderedCount = (uint32_t*)(d + oCnt);
    p.back         = (uint32_t*)(d + oBack);
    p.pathEmission = (uint32_t*)(d + oPe);
    p.pathState    = pathState ? (uint32_t*)(d + oPs) : nullptr;
    p.pathScore    = (float*)(d + oSc);
    if (smemBytes > 48 * 1024)
        GMM_CUDA_TRY(h, cudaFuncSetAttribute(align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
    align_kernel<<<grid, 256, smemBytes, h->stream>>>(p);
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    GMM_CUDA_TRY(h, cudaMemcpyAsync(pathEmission, d + oPe, (size_t)nFrames * 4, cudaMemcpyDeviceToHost, h->stream));
    if (pathState)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(pathState, d + oPs, (size_t)nFrames * 4, cudaMemcpyDeviceToHost, h->stream));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(pathScore, d + oSc, (size_t)nSeg * 4, cudaMemcpyDeviceToHost, h->stream));
    uint32_t ordered = 0;
    GMM_CUDA_TRY(h, cudaMemcpyAsync(&ordered, d + oCnt, 4, cudaMemcpyDeviceToHost, h->stream));
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    h->alignOrdered += ordered;
    return GMM_OK;
}
