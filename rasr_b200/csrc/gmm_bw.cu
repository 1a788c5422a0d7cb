// Baum-Welch alignment on the device (second mode of the reference's aligner): forward-backward
// over alignment automata, reading the emission scores where the scoring kernels left them; what leaves the kernel is,
// per frame, the posterior of every emission the segment's automaton names -- the weights of the reference's
// alignment items -- in a block laid out like the compact score block (column j = entry j of the emission list).
//
// Replaces, for unpruned Baum-Welch alignment with min-weight 0 (the defaults):
//   Search::Aligner::SearchSpace::activateOrUpdateStateHypothesisBaumWelch / expand   Search/Aligner.cc:160-187,190-239
//       forward potential of a state = log-semiring sum over its incoming arcs of
//       potential(source) + (transition weight + emission score)
//   getAlignmentFsaBaumWelch + Aligner::getAlignmentPosteriorFsa (Fsa::posterior64)   Aligner.cc:438-452, 757-772
//       arc posteriors from forward and backward potentials
//   Aligner::getAlignment + Speech::Alignment::combineItems / clipWeights / expm / normalizeWeights
//       Aligner.cc:693-716, Speech/Alignment.cc:455-566: posteriors summed per (frame, emission), clipped to <= 1
// Not taken over: acoustic pruning (Aligner.cc:241-330), min-weight > 0 (Fsa::prunePosterior re-weights the lattice).
//
// Numerics.  The reference computes posteriors in f64 over f32 arc scores (Fsa::posterior64).  Here the potentials are
// f64 as well (shared memory and the stored forward pass), every log-sum-exp subtracts its minimum in f64 and
// evaluates expf / logf on the small f32 remainders, so nothing rounds at the magnitude of a path score (3e5 after
// 4000 frames, where f32 has an ulp of 0.03).
//
// One CTA per segment, one thread per state (strided when the automaton is larger than the block).  Forward pass:
// incoming-arc lists (pull), relative potentials of every frame stored to global memory.  Backward pass: the caller's
// own (outgoing) arc lists -- a source state pulls beta from its targets and, in the same loop, adds the arc
// posteriors into the frame's emission row in shared memory (f32 atomics), which is then written out as one row.
// Emission scores and stored potentials of the next frames arrive through a ring of asynchronous copies.
#include <float.h>
#include <math.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <vector>

#include "gmm_internal.cuh"

namespace {

constexpr int kRingFrames = 4;

struct BwSeg {
    uint32_t frame0, frames, state0, states, initial, arc0, arcs, em0, nEm;
    uint32_t pad;
    uint64_t alpha0;  // offset of this segment's stored potentials [(frames + 1) x states]
    uint64_t pad2;
};

struct BwParams {
    const BwSeg*    seg;
    uint32_t        nSeg;
    const float*    scores;
    size_t          ld;
    uint32_t        compact;
    const uint32_t* emList;
    const uint32_t *inOff, *inSrc, *inSlot;  // incoming arcs (CSR by global state; positions global)
    const float*    inW;
    const uint32_t *outOff, *outTgt, *outSlot;  // the caller's arc lists
    const float *   outW, *finalW;
    double*         alpha;
    float*          weights;  // [frames x ldW]
    size_t          ldW;
    double*         total;    // [nSeg]
};

__device__ __forceinline__ void cp4(void* smemDst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smemDst)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp8(void* smemDst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smemDst)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_commit() {
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template<int N>
__device__ __forceinline__ void cp_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
        v = fminf(v, __shfl_xor_sync(0xFFFFFFFFu, v, d));
    return v;
}

size_t bw_smem_words(uint32_t S, uint32_t arcs, uint32_t nEm) {  // in 4-byte words
    return 2 * ((2 + (size_t)kRingFrames) * S + 32) + 2 * ((size_t)S + 1) + 6 * (size_t)arcs + (size_t)kRingFrames * nEm +
           nEm + 32 + 8;
}

__global__ void __launch_bounds__(1024) bw_kernel(const BwParams p) {
    extern __shared__ double smd[];
    constexpr int  K   = kRingFrames;
    const uint32_t tid = threadIdx.x, nThreads = blockDim.x, warp = tid >> 5, lane = tid & 31, nWarps = nThreads >> 5;
    for (uint32_t si = blockIdx.x; si < p.nSeg; si += gridDim.x) {
        const BwSeg    sg = p.seg[si];
        const uint32_t S = sg.states, T = sg.frames, A = sg.arcs, nEm = sg.nEm;
        double*        va     = smd;                             // potentials, ping
        double*        vb     = va + S;                          // pong
        double*        ringA  = vb + S;                          // [K][S] stored forward potentials (backward pass)
        double*        redD   = ringA + (size_t)K * S;           // [32] reductions
        uint32_t*      inOff  = (uint32_t*)(redD + 32);
        uint32_t*      outOff = inOff + S + 1;
        uint32_t*      inSrc  = outOff + S + 1;
        float*         inW    = (float*)(inSrc + A);
        uint32_t*      inSlot = (uint32_t*)(inW + A);
        uint32_t*      outTgt = inSlot + A;
        float*         outW   = (float*)(outTgt + A);
        uint32_t*      outSl  = (uint32_t*)(outW + A);
        float*         ringE  = (float*)(outSl + A);             // [K][nEm] emission scores of the frames in flight
        float*         gam    = ringE + (size_t)K * nEm;         // [nEm] posteriors of the current frame
        float*         red    = gam + nEm;                       // [32]
        double*        alphaG = p.alpha + sg.alpha0;
        float*         wOut   = p.weights + (size_t)sg.frame0 * p.ldW;
        for (uint32_t s = tid; s <= S; s += nThreads) {
            inOff[s]  = p.inOff[sg.state0 + s] - sg.arc0;
            outOff[s] = p.outOff[sg.state0 + s] - sg.arc0;
        }
        for (uint32_t a = tid; a < A; a += nThreads) {
            inSrc[a]  = p.inSrc[sg.arc0 + a];
            inW[a]    = p.inW[sg.arc0 + a];
            inSlot[a] = p.inSlot[sg.arc0 + a];
            outTgt[a] = p.outTgt[sg.arc0 + a];
            outW[a]   = p.outW[sg.arc0 + a];
            outSl[a]  = p.outSlot[sg.arc0 + a];
        }
        for (uint32_t s = tid; s < S; s += nThreads) {
            va[s]     = (s == sg.initial) ? 0.0 : (double)INFINITY;
            alphaG[s] = va[s];
        }
        for (uint32_t j = tid; j < nEm; j += nThreads)
            gam[j] = 0.f;
        __syncthreads();
        auto fetchE = [&](uint32_t f) {  // emission scores of frame f into its ring slot
            const float* row = p.scores + (size_t)(sg.frame0 + f) * p.ld;
            float*       dst = ringE + (size_t)(f % K) * nEm;
            for (uint32_t e = tid; e < nEm; e += nThreads)
                cp4(dst + e, row + (p.compact ? e : p.emList[sg.em0 + e]));
        };
        // ---- forward: alpha[t + 1][s] = -log sum_in exp(-(alpha[t][src] + (w + e))) ----------------------------------
        for (uint32_t f = 0; f + 1 < (uint32_t)K; ++f) {
            if (f < T)
                fetchE(f);
            cp_commit();
        }
        double* cur = va;
        double* nxt = vb;
        for (uint32_t t = 0; t < T; ++t) {
            cp_wait<K - 2>();
            __syncthreads();
            if (t + K - 1 < T)
                fetchE(t + K - 1);
            cp_commit();
            const float* E = ringE + (size_t)(t % K) * nEm;
            for (uint32_t s = tid; s < S; s += nThreads) {
                const uint32_t a0 = inOff[s], a1 = inOff[s + 1];
                double         mn = INFINITY;
                for (uint32_t a = a0; a < a1; ++a)
                    mn = fmin(mn, cur[inSrc[a]] + (double)(inW[a] + E[inSlot[a]]));
                double v = INFINITY;
                if (mn < (double)INFINITY) {
                    float sum = 0.f;
                    for (uint32_t a = a0; a < a1; ++a)
                        sum += expf((float)(mn - (cur[inSrc[a]] + (double)(inW[a] + E[inSlot[a]]))));
                    v = mn - (double)logf(sum);
                }
                nxt[s]                           = v;
                alphaG[(size_t)(t + 1) * S + s] = v;
            }
            double* sw = cur;
            cur        = nxt;
            nxt        = sw;
        }
        cp_wait<0>();
        __syncthreads();
        // ---- total = -log sum_final exp(-(alpha[T][s] + finalWeight[s])) -------------------------------------------
        double fmn = INFINITY;
        for (uint32_t s = tid; s < S; s += nThreads) {
            const float fw = p.finalW[sg.state0 + s];
            if (fw < INFINITY)
                fmn = fmin(fmn, cur[s] + (double)fw);
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
            fmn = fmin(fmn, __shfl_xor_sync(0xFFFFFFFFu, fmn, d));
        if (lane == 0)
            redD[warp] = fmn;
        __syncthreads();
        fmn = INFINITY;
        for (uint32_t w = 0; w < nWarps; ++w)
            fmn = fmin(fmn, redD[w]);
        const bool reached = fmn < (double)INFINITY;
        float      fsum    = 0.f;
        if (reached)
            for (uint32_t s = tid; s < S; s += nThreads) {
                const float fw = p.finalW[sg.state0 + s];
                if (fw < INFINITY)
                    fsum += expf((float)(fmn - (cur[s] + (double)fw)));
            }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
            fsum += __shfl_xor_sync(0xFFFFFFFFu, fsum, d);
        if (lane == 0)
            red[warp] = fsum;
        __syncthreads();
        fsum = 0.f;
        for (uint32_t w = 0; w < nWarps; ++w)
            fsum += red[w];
        const double total = reached ? fmn - (double)logf(fsum) : (double)FLT_MAX;
        if (tid == 0)
            p.total[si] = total;
        __syncthreads();
        if (!reached) {  // empty alignment (Aligner.cc:734-735): no items
            for (uint32_t t = 0; t < T; ++t)
                for (uint32_t j = tid; j < nEm; j += nThreads)
                    wOut[(size_t)t * p.ldW + j] = 0.f;
            __syncthreads();
            continue;
        }
        // ---- backward: beta[t][src] = -log sum_out exp(-((w + e) + beta[t + 1][tgt])), arc posteriors ---------------
        auto fetchB = [&](uint32_t f) {  // frame f: its emission scores and the potentials BEFORE it
            fetchE(f);
            double* dst = ringA + (size_t)(f % K) * S;
            for (uint32_t s = tid; s < S; s += nThreads)
                cp8(dst + s, alphaG + (size_t)f * S + s);
        };
        for (uint32_t k = 0; k + 1 < (uint32_t)K; ++k) {
            if (k < T)
                fetchB(T - 1 - k);
            cp_commit();
        }
        double* bn = nxt;  // beta[t + 1]
        double* bc = cur;  // beta[t]   (cur = alpha[T] is no longer needed: it is in global memory)
        for (uint32_t s = tid; s < S; s += nThreads)
            bn[s] = (double)p.finalW[sg.state0 + s];
        for (uint32_t t = T; t-- > 0;) {
            cp_wait<K - 2>();
            __syncthreads();
            if (t >= (uint32_t)(K - 1))
                fetchB(t - (K - 1));
            cp_commit();
            const float*  E  = ringE + (size_t)(t % K) * nEm;
            const double* Al = ringA + (size_t)(t % K) * S;
            for (uint32_t s = tid; s < S; s += nThreads) {
                const uint32_t a0 = outOff[s], a1 = outOff[s + 1];
                double         mn = INFINITY;
                for (uint32_t a = a0; a < a1; ++a)
                    mn = fmin(mn, (double)(outW[a] + E[outSl[a]]) + bn[outTgt[a]]);
                double v = INFINITY;
                if (mn < (double)INFINITY) {
                    const double al  = Al[s] - total;
                    float        sum = 0.f;
                    for (uint32_t a = a0; a < a1; ++a) {
                        const double r = (double)(outW[a] + E[outSl[a]]) + bn[outTgt[a]];
                        sum += expf((float)(mn - r));
                        if (al < (double)INFINITY && r < (double)INFINITY)
                            atomicAdd(&gam[outSl[a]], expf(-(float)(al + r)));  // the arc's posterior
                    }
                    v = mn - (double)logf(sum);
                }
                bc[s] = v;
            }
            __syncthreads();
            // clipWeights (a posterior is at most 1), then normalizeWeights over the frame (Alignment.cc:530-547)
            float part = 0.f;
            for (uint32_t j = tid; j < nEm; j += nThreads)
                part += fminf(gam[j], 1.f);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1)
                part += __shfl_xor_sync(0xFFFFFFFFu, part, d);
            if (lane == 0)
                red[warp] = part;
            __syncthreads();
            float frameSum = 0.f;
            for (uint32_t w = 0; w < nWarps; ++w)
                frameSum += red[w];
            const float inv = frameSum > 0.f ? 1.f / frameSum : 1.f;
            for (uint32_t j = tid; j < nEm; j += nThreads) {
                wOut[(size_t)t * p.ldW + j] = fminf(gam[j], 1.f) * inv;
                gam[j]                      = 0.f;
            }
            double* sw = bn;
            bn         = bc;
            bc         = sw;
        }
        cp_wait<0>();
        __syncthreads();
    }
}

}  // namespace

// Host arrays in; `d_scores` is the DEVICE score block (full: column = emission index, or compact: column j = entry j
// of the segment's emission list), `d_weights` the DEVICE posterior block [frames x ldW] (ldW >= longest emission
// list), totalScore (host, may be NULL) -log p(x | automaton) per segment, FLT_MAX for an unreachable one.
int gmm_bw_launch(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                  const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                  const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                  const float* finalWeight, float* d_weights, size_t ldW, double* totalScore, bool compact,
                  uint32_t nEmissions, const AlignLists* lists, bool wait, BwDevice* dev) {
    if (nSeg == 0)
        return GMM_OK;
    const uint32_t    nStates = stateOffsets[nSeg], nFrames = frameOffsets[nSeg], nArcs = arcOffsets[nStates];
    AlignLists        own;
    const AlignLists* L       = lists;
    const uint32_t    emLimit = compact ? nEmissions : (uint32_t)std::min<size_t>(ld, 0xFFFFFFFFu);
    if (!L) {
        if (!gmm_align_emission_lists(h, nSeg, stateOffsets, arcOffsets, arcEmission, emLimit, own)) {
            h->setError("gmm_baum_welch: arc offsets not ascending or emission index outside the score block");
            return GMM_ERR_ARG;
        }
        L = &own;
    }
    auto   al   = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t off  = 0;
    auto   take = [&](size_t bytes) {
        size_t o = off;
        off += al(bytes);
        return o;
    };
    const size_t oSeg = take(nSeg * sizeof(BwSeg)), oInOff = take((nStates + 1) * 4), oInSrc = take((size_t)nArcs * 4),
                 oInW = take((size_t)nArcs * 4), oInSlot = take((size_t)nArcs * 4), oOutOff = take((nStates + 1) * 4),
                 oOutTgt = take((size_t)nArcs * 4), oOutW = take((size_t)nArcs * 4), oOutSlot = take((size_t)nArcs * 4),
                 oFin = take((size_t)nStates * 4), oEmList = take(L->emList.size() * 4), oFrameOff = take((nSeg + 1) * 4),
                 oEmOff = take((nSeg + 1) * 4), tableBytes = off;
    void* pin = nullptr;
    int   rc  = h->reservePinned(PIN_ALIGN, tableBytes, &pin);
    if (rc)
        return rc;
    char*     hp    = (char*)pin;
    BwSeg*    segs  = (BwSeg*)(hp + oSeg);
    uint32_t *inOff = (uint32_t*)(hp + oInOff), *inSrc = (uint32_t*)(hp + oInSrc), *inSlot = (uint32_t*)(hp + oInSlot);
    float*    inW   = (float*)(hp + oInW);
    uint64_t  alphaTotal = 0, offTotal = 0;
    uint32_t  maxStates = 0, maxList = 0;
    size_t    maxWords = 0;
    for (uint32_t i = 0; i < nSeg; ++i) {
        const uint32_t s0 = stateOffsets[i], s1 = stateOffsets[i + 1];
        if (s1 < s0 || frameOffsets[i + 1] < frameOffsets[i] || initialState[i] >= std::max(1u, s1 - s0)) {
            h->setError("gmm_baum_welch: offsets not ascending or initial state out of range");
            return GMM_ERR_ARG;
        }
        BwSeg& sg  = segs[i];
        sg.frame0  = frameOffsets[i];
        sg.frames  = frameOffsets[i + 1] - frameOffsets[i];
        sg.state0  = s0;
        sg.states  = s1 - s0;
        sg.initial = initialState[i];
        sg.arc0    = arcOffsets[s0];
        sg.arcs    = arcOffsets[s1] - arcOffsets[s0];
        sg.em0     = L->emOff[i];
        sg.nEm     = L->emOff[i + 1] - L->emOff[i];
        sg.pad     = 0;
        sg.alpha0  = alphaTotal;
        sg.pad2    = 0;
        alphaTotal += (uint64_t)(sg.frames + 1) * sg.states;
        offTotal += sg.frames + 1;
        maxStates = std::max(maxStates, sg.states);
        maxList   = std::max(maxList, sg.nEm);
        maxWords  = std::max(maxWords, bw_smem_words(sg.states, sg.arcs, sg.nEm));
    }
    if ((size_t)maxList > ldW || (compact && (size_t)maxList > ld)) {
        h->setError("gmm_baum_welch: posterior (or compact score) block narrower than a segment's emission list");
        return GMM_ERR_ARG;
    }
    const size_t smemBytes = maxWords * 4 + 16;
    if (smemBytes > 200 * 1024) {
        h->setError("gmm_baum_welch: automaton too large for the kernel's shared-memory tables");
        return GMM_ERR_UNSUPPORTED;
    }
    // incoming-arc lists per segment (targets are segment-local, so a segment's lists occupy its own arc range)
    HostPool&        pool = h->hostPool();
    std::atomic<int> bad(0);
    inOff[0] = arcOffsets[0];
    pool.run(nSeg, std::max(1u, nSeg / (4 * (uint32_t)pool.size())), [&](uint32_t i0, uint32_t i1, int worker) {
        std::vector<uint32_t>& fill = h->workerWords[worker];
        if (fill.size() < (size_t)maxStates + 1)
            fill.resize((size_t)maxStates + 1);
        for (uint32_t i = i0; i < i1; ++i) {
            const uint32_t st0 = stateOffsets[i], S = stateOffsets[i + 1] - st0;
            std::fill(fill.begin(), fill.begin() + S + 1, 0u);
            for (uint32_t s = st0; s < st0 + S; ++s) {
                if (arcOffsets[s + 1] < arcOffsets[s]) {
                    bad = 1;
                    return;
                }
                for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a) {
                    if (arcTarget[a] >= S) {
                        bad = 2;
                        return;
                    }
                    fill[arcTarget[a] + 1]++;
                }
            }
            uint32_t run = arcOffsets[st0];
            for (uint32_t s = 0; s < S; ++s) {
                const uint32_t c   = fill[s + 1];
                fill[s]            = run;
                run += c;
                inOff[st0 + s + 1] = run;
            }
            for (uint32_t s = st0; s < st0 + S; ++s)
                for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a) {
                    const uint32_t k = fill[arcTarget[a]]++;
                    inSrc[k]         = s - st0;
                    inW[k]           = arcWeight[a];
                    inSlot[k]        = L->arcSlot[a];
                }
        }
    });
    if (bad) {
        h->setError(bad == 1 ? "gmm_baum_welch: arc offsets not ascending" : "gmm_baum_welch: arc target outside its segment's automaton");
        return GMM_ERR_ARG;
    }
    memcpy(hp + oOutOff, arcOffsets, (size_t)(nStates + 1) * 4);
    if (nArcs) {
        memcpy(hp + oOutTgt, arcTarget, (size_t)nArcs * 4);
        memcpy(hp + oOutW, arcWeight, (size_t)nArcs * 4);
        memcpy(hp + oOutSlot, L->arcSlot.data(), (size_t)nArcs * 4);
    }
    memcpy(hp + oFin, finalWeight, (size_t)nStates * 4);
    if (!L->emList.empty())
        memcpy(hp + oEmList, L->emList.data(), L->emList.size() * 4);
    memcpy(hp + oFrameOff, frameOffsets, (size_t)(nSeg + 1) * 4);
    memcpy(hp + oEmOff, L->emOff.data(), (size_t)(nSeg + 1) * 4);
    const size_t oTotal = take((size_t)nSeg * 8), oAlpha = take((size_t)alphaTotal * 8);
    (void)offTotal;
    void*        base   = nullptr;
    if ((rc = h->reserve(SCR_MISC, off, &base)))
        return rc;
    char* d = (char*)base;
    if ((rc = gmm_copy_tables(h, d, hp, tableBytes)))
        return rc;
    if ((rc = h->pinnedCopied(PIN_ALIGN)))
        return rc;
    BwParams p;
    p.seg     = (const BwSeg*)(d + oSeg);
    p.nSeg    = nSeg;
    p.scores  = d_scores;
    p.ld      = ld;
    p.compact = compact ? 1u : 0u;
    p.emList  = (const uint32_t*)(d + oEmList);
    p.inOff   = (const uint32_t*)(d + oInOff);
    p.inSrc   = (const uint32_t*)(d + oInSrc);
    p.inSlot  = (const uint32_t*)(d + oInSlot);
    p.inW     = (const float*)(d + oInW);
    p.outOff  = (const uint32_t*)(d + oOutOff);
    p.outTgt  = (const uint32_t*)(d + oOutTgt);
    p.outSlot = (const uint32_t*)(d + oOutSlot);
    p.outW    = (const float*)(d + oOutW);
    p.finalW  = (const float*)(d + oFin);
    p.alpha   = (double*)(d + oAlpha);
    p.weights = d_weights;
    p.ldW     = ldW;
    p.total   = (double*)(d + oTotal);
    if (smemBytes > 48 * 1024)
        GMM_CUDA_TRY(h, cudaFuncSetAttribute(bw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
    const int grid    = (int)std::min<uint32_t>(nSeg, (uint32_t)h->smCount * 8);
    const int threads = (int)std::min<uint32_t>(1024, std::max<uint32_t>(128, (maxStates + 31) / 32 * 32));
    (void)nFrames;
    h->timeBegin();
    bw_kernel<<<grid, threads, smemBytes, h->stream>>>(p);
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    if (dev) {
        dev->frameOff = (const uint32_t*)(d + oFrameOff);
        dev->emOff    = (const uint32_t*)(d + oEmOff);
        dev->emList   = (const uint32_t*)(d + oEmList);
        dev->total    = (const double*)(d + oTotal);
    }
    if (totalScore)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(totalScore, d + oTotal, (size_t)nSeg * 8, cudaMemcpyDeviceToHost, h->stream));
    if (wait || totalScore)
        GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GMM_OK;
}
