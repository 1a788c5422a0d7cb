// Mixture-sharded scoring with its exchange step: every rank scores its own mixtures into its own frame-major block
// [T][ld] and PUSHES those columns into the peers' blocks (mapped through CUDA IPC) with full-line stores over
// NVLink, one column range at a time while the tensor kernel is already scoring the next range.  The ranks
// synchronise on the device through flag words in the blocks themselves; there is no host barrier and no NCCL call.
//
// Replaces the all-gather of per-frame score columns of BASELINE.json configs[4] (the reference has no multi-GPU
// scorer: Mm::FeatureScorer::Scorer::score(e) is answered from one process' own mixture set,
// src/Mm/FeatureScorer.hh:71-116).
//
// Flag words of a block (32 x u32 behind the scores, zeroed by gmm_peer_alloc):
//   [r]      READY from rank r: r has finished reading the previous result in r's own block, which may now be
//            overwritten (written by r into every peer's flags; the push kernel waits for it before it writes to r)
//   [8 + r]  DONE from rank r: all of r's columns of this epoch have landed in this block
//   [16]     CTA counter of the push kernel, [17] time-out marker
#include "gmm_internal.cuh"

#include <algorithm>
#include <type_traits>
#include <cstdlib>

namespace {

constexpr uint32_t kFlagReady = 0, kFlagDone = 8, kFlagCount = 16, kFlagTimeout = 17;
constexpr long long kSpinLimit = 4000000000ll;  // ~2 s of SM clocks: a missing peer ends the wait instead of hanging

struct GatherPeers {
    float*    block[8];
    uint32_t* flags[8];
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// epochs wrap: "reached" = not behind
__device__ __forceinline__ bool reached(uint32_t have, uint32_t want) { return (int32_t)(have - want) >= 0; }

__device__ bool spin_until(const uint32_t* flag, uint32_t epoch, uint32_t* timeoutMark) {
    const long long t0 = clock64();
    while (!reached(ld_acquire_sys(flag), epoch)) {
        if (clock64() - t0 > kSpinLimit) {
            *timeoutMark = 1;
            return false;
        }
        __nanosleep(64);
    }
    return true;
}

// Everything this rank launched before on its stream has finished reading its block: tell the peers they may write
// this epoch's columns into it.
__global__ void gather_signal_ready(GatherPeers peers, uint32_t rank, uint32_t world, uint32_t epoch) {
    const uint32_t r = threadIdx.x;
    if (r < world && r != rank)
        st_release_sys(peers.flags[r] + kFlagReady + rank, epoch);
}

// Copy columns [col0, col1) of rows [0, T) of the own block into the same place of every peer's block.
// VEC = floats per access (4 when the range and ld are 16-byte aligned).  A thread reads kUnroll accesses (the own
// block is an L2 hit for short calls, HBM for long ones: several loads in flight per thread) and stores each of them
// to all peers in turn, so the source is read once and consecutive stores leave on different links.
template<int VEC, int kUnroll, typename IDX>
__device__ __forceinline__ void push_range(const GatherPeers& peers, uint32_t rank, uint32_t world, IDX total, uint32_t w,
                                           size_t ld, uint32_t col0) {
    typedef typename std::conditional<VEC == 4, float4, float>::type V;
    const float* src    = peers.block[rank];
    const IDX    stride = (IDX)gridDim.x * blockDim.x;
    for (IDX i0 = (IDX)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += stride * kUnroll) {
        V      v[kUnroll];
        size_t off[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const IDX i = i0 + (IDX)u * stride;
            if (i < total) {
                const IDX row = i / w, c = i - row * w;
                off[u]        = (size_t)row * ld + col0 + (size_t)c * VEC;
                v[u]          = *reinterpret_cast<const V*>(src + off[u]);
            }
        }
        for (uint32_t k = 1; k < world; ++k) {
            uint32_t r = rank + k;
            r          = r >= world ? r - world : r;
            float* dst = peers.block[r];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u)
                if (i0 + (IDX)u * stride < total)
                    *reinterpret_cast<V*>(dst + off[u]) = v[u];
        }
    }
}

template<int VEC, int kUnroll>
__global__ void __launch_bounds__(256) gather_push_kernel(GatherPeers peers, uint32_t rank, uint32_t world, uint32_t T,
                                                          size_t ld, uint32_t col0, uint32_t col1, uint32_t epoch,
                                                          int last) {
    uint32_t* own = peers.flags[rank];
    __shared__ int ok;
    if (threadIdx.x == 0)
        ok = 1;
    __syncthreads();
    if (threadIdx.x < world && threadIdx.x != rank)
        if (!spin_until(own + kFlagReady + threadIdx.x, epoch, own + kFlagTimeout))
            ok = 0;
    __syncthreads();
    if (ok) {
        const uint32_t w     = (col1 - col0) / VEC;  // accesses per row
        const size_t   total = (size_t)T * w;
        if (total + (size_t)gridDim.x * blockDim.x * kUnroll <= 0xFFFFFFFFull)
            push_range<VEC, kUnroll, uint32_t>(peers, rank, world, (uint32_t)total, w, ld, col0);
        else
            push_range<VEC, kUnroll, size_t>(peers, rank, world, total, w, ld, col0);
    }
    if (!last)
        return;
    // the last CTA to finish tells every peer that this rank's columns are complete
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t before = atomicAdd(own + kFlagCount, 1u);
        ok                    = before + 1 == gridDim.x;
    }
    __syncthreads();
    if (ok) {
        __threadfence_system();
        if (threadIdx.x == 0)
            own[kFlagCount] = 0;
        if (threadIdx.x < world && threadIdx.x != rank)
            st_release_sys(peers.flags[threadIdx.x] + kFlagDone + rank, epoch);
    }
}

// Every peer's columns of this epoch have landed in the own block.
__global__ void gather_wait_done(GatherPeers peers, uint32_t rank, uint32_t world, uint32_t epoch) {
    uint32_t* own = peers.flags[rank];
    if (threadIdx.x < world && threadIdx.x != rank)
        spin_until(own + kFlagDone + threadIdx.x, epoch, own + kFlagTimeout);
}

}  // namespace

int gmm_gather_score(gmm_handle* h, uint32_t T, const float* d_x, uint32_t rank, uint32_t world,
                     void* const* d_blocks, size_t ld, uint32_t colOffset, size_t flagsOffsetBytes, uint32_t chunks) {
    const TensorModel& tm = h->tensor;
    GatherPeers        peers = {};
    for (uint32_t r = 0; r < world; ++r) {
        peers.block[r] = reinterpret_cast<float*>(d_blocks[r]);
        peers.flags[r] = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(d_blocks[r]) + flagsOffsetBytes);
    }
    if (!h->gatherStream) {
        GMM_CUDA_TRY(h, cudaStreamCreateWithFlags(&h->gatherStream, cudaStreamNonBlocking));
        for (int i = 0; i < gmm_handle::kGatherEvents; ++i)
            GMM_CUDA_TRY(h, cudaEventCreateWithFlags(&h->gatherEvent[i], cudaEventDisableTiming));
    }
    const uint32_t epoch = ++h->gatherEpoch;
    float*         own   = peers.block[rank] + colOffset;
    // column ranges follow whole B tiles; only a regular model has its tiles at known columns
    if (!tm.regular || chunks == 0)
        chunks = 1;
    chunks = std::min<uint32_t>(std::min<uint32_t>(chunks, gmm_handle::kGatherEvents - 1), std::max(1u, tm.nTiles));
    static const int pushCtas = getenv("GMM_GATHER_CTAS") ? std::max(1, atoi(getenv("GMM_GATHER_CTAS"))) : 128;

    static const int unrollEnv = getenv("GMM_GATHER_UNROLL") ? atoi(getenv("GMM_GATHER_UNROLL")) : 0;
    // measured at N=2 (profiles/sweep_push_n2_*_r01.txt): one load in flight per thread and 128 CTAs disturb the tensor
    // kernel running beside the push least (T=8192, 8 ranges: 0.951 ms; four loads in flight: 1.065 ms)
    const int        unroll    = unrollEnv ? unrollEnv : 1;

    gather_signal_ready<<<1, 32, 0, h->stream>>>(peers, rank, world, epoch);
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    for (uint32_t c = 0; c < chunks; ++c) {
        const uint32_t t0 = (uint32_t)((uint64_t)tm.nTiles * c / chunks), t1 = (uint32_t)((uint64_t)tm.nTiles * (c + 1) / chunks);
        const int      rc = gmm_tc_score_tiles(h, T, d_x, own, ld, t0, t1, c == 0);
        if (rc)
            return rc;
        const uint32_t col0 = colOffset + (chunks == 1 ? 0u : t0 * tm.mixPerTile);
        const uint32_t col1 = colOffset + (chunks == 1 || t1 == tm.nTiles ? h->M : t1 * tm.mixPerTile);
        if (world > 1) {
            GMM_CUDA_TRY(h, cudaEventRecord(h->gatherEvent[c], h->stream));
            GMM_CUDA_TRY(h, cudaStreamWaitEvent(h->gatherStream, h->gatherEvent[c], 0));
            bool vec = ld % 4 == 0 && col0 % 4 == 0 && (col1 - col0) % 4 == 0;
            for (uint32_t r = 0; r < world && vec; ++r)
                vec = reinterpret_cast<uintptr_t>(peers.block[r]) % 16 == 0;
            const int last = c + 1 == chunks;
#define GMM_PUSH(V, U) \
    gather_push_kernel<V, U><<<pushCtas, 256, 0, h->gatherStream>>>(peers, rank, world, T, ld, col0, col1, epoch, last)
            if (vec) {
                if (unroll == 1)
                    GMM_PUSH(4, 1);
                else if (unroll == 2)
                    GMM_PUSH(4, 2);
                else
                    GMM_PUSH(4, 4);
            }
            else
                GMM_PUSH(1, 1);
#undef GMM_PUSH
            h->launches++;
            GMM_CUDA_TRY(h, cudaGetLastError());
        }
    }
    if (world > 1) {
        cudaEvent_t done = h->gatherEvent[gmm_handle::kGatherEvents - 1];
        GMM_CUDA_TRY(h, cudaEventRecord(done, h->gatherStream));
        GMM_CUDA_TRY(h, cudaStreamWaitEvent(h->stream, done, 0));
        gather_wait_done<<<1, 32, 0, h->stream>>>(peers, rank, world, epoch);
        h->launches++;
        GMM_CUDA_TRY(h, cudaGetLastError());
    }
    return GMM_OK;
}
