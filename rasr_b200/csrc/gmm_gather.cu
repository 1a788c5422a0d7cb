// Mixture-sharded scoring with its exchange step: every rank scores its own mixtures into its own frame-major block
// [T][ld] and PUSHES those columns into the peers' blocks (mapped through CUDA IPC) with full-line stores over
// NVLink, one range (of columns or of frames) at a time while the tensor kernel is already scoring the next.  The ranks
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
//   [16]     CTA counter of the push kernel, [17] time-out marker, [18] epoch of the last call this rank made on
//            the block (the epoch lives WITH the flag words: a new handle over a used block continues its sequence)
// A time-out is not silent: the waiting kernel stores the epoch into a page-locked word the host can read
// (gmm_gather_status), fills every column of the call's rows that is not this rank's own with +inf -- a consumer already
// queued behind the call sees impossible scores, never stale ones -- and the next call on the handle that scores or
// consumes scores returns GMM_ERR_TIMEOUT until gmm_gather_reset.
#include "gmm_internal.cuh"

#include <algorithm>
#include <type_traits>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace {

constexpr uint32_t kFlagReady = 0, kFlagDone = 8, kFlagCount = 16, kFlagTimeout = 17, kFlagEpoch = 18, kFlagWords = 32;
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

__device__ bool spin_until(const uint32_t* flag, uint32_t epoch, uint32_t* timeoutMark, uint32_t* hostStatus,
                           long long limit) {
    const long long t0 = clock64();
    while (!reached(ld_acquire_sys(flag), epoch)) {
        if (clock64() - t0 > limit) {
            *timeoutMark = 1;
            *reinterpret_cast<volatile uint32_t*>(hostStatus) = epoch ? epoch : 1u;  // page-locked host word
            __threadfence_system();
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

// Copy columns [col0, col1) of rows [row0, row0 + T) of the own block into the same place of every peer's block.
// VEC = floats per access (4 when the range and ld are 16-byte aligned).  A thread reads kUnroll accesses (the own
// block is an L2 hit for short calls, HBM for long ones: several loads in flight per thread) and stores each of them
// to all peers in turn, so the source is read once and consecutive stores leave on different links.
template<int VEC, int kUnroll, typename IDX>
__device__ __forceinline__ void push_range(const GatherPeers& peers, uint32_t rank, uint32_t world, IDX total, uint32_t w,
                                           size_t ld, uint32_t col0, uint32_t row0) {
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
                off[u]        = ((size_t)row0 + row) * ld + col0 + (size_t)c * VEC;
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
__global__ void __launch_bounds__(256) gather_push_kernel(GatherPeers peers, uint32_t rank, uint32_t world, uint32_t row0,
                                                          uint32_t T, size_t ld, uint32_t col0, uint32_t col1,
                                                          uint32_t epoch, int last, uint32_t* hostStatus,
                                                          long long limit) {
    uint32_t* own = peers.flags[rank];
    __shared__ int ok;
    if (threadIdx.x == 0)
        ok = 1;
    __syncthreads();
    if (threadIdx.x < world && threadIdx.x != rank)
        if (!spin_until(own + kFlagReady + threadIdx.x, epoch, own + kFlagTimeout, hostStatus, limit))
            ok = 0;
    __syncthreads();
    if (ok) {
        const uint32_t w     = (col1 - col0) / VEC;  // accesses per row
        const size_t   total = (size_t)T * w;
        if (total + (size_t)gridDim.x * blockDim.x * kUnroll <= 0xFFFFFFFFull)
            push_range<VEC, kUnroll, uint32_t>(peers, rank, world, (uint32_t)total, w, ld, col0, row0);
        else
            push_range<VEC, kUnroll, size_t>(peers, rank, world, total, w, ld, col0, row0);
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
        // after a time-out (this or an earlier range of the call) the columns are NOT complete: say nothing, the
        // peers then run into their own time-out instead of reading stale scores
        const bool complete = *reinterpret_cast<volatile uint32_t*>(own + kFlagTimeout) == 0;
        if (complete && threadIdx.x < world && threadIdx.x != rank)
            st_release_sys(peers.flags[threadIdx.x] + kFlagDone + rank, epoch);
    }
}

// Copy-engine variant of the push (long ranges): the READY wait and the DONE announcement as one-warp kernels around
// cudaMemcpy2DAsync calls on the push stream.  The copy engines move a range at the NVLink rate without taking
// registers or issue slots from the scoring kernel, which leaves a co-resident push kernel one thin CTA per SM
// (measured: 275 GB/s leaving the GPU under the scoring kernel against 650-700 GB/s alone).
__global__ void gather_wait_ready(GatherPeers peers, uint32_t rank, uint32_t world, uint32_t epoch, uint32_t* hostStatus,
                                  long long limit) {
    uint32_t* own = peers.flags[rank];
    if (threadIdx.x < world && threadIdx.x != rank)
        spin_until(own + kFlagReady + threadIdx.x, epoch, own + kFlagTimeout, hostStatus, limit);
}
__global__ void gather_announce_done(GatherPeers peers, uint32_t rank, uint32_t world, uint32_t epoch) {
    uint32_t* own = peers.flags[rank];
    __threadfence_system();
    // after a time-out the columns are not complete: say nothing (the peers run into their own time-out)
    const bool complete = *reinterpret_cast<volatile uint32_t*>(own + kFlagTimeout) == 0;
    if (complete && threadIdx.x < world && threadIdx.x != rank)
        st_release_sys(peers.flags[threadIdx.x] + kFlagDone + rank, epoch);
}

// Every peer's columns of this epoch have landed in the own block -- or, after a time-out here or in one of the call's
// push kernels, none of them counts: rows [0, T) keep the own columns [col0, col1) and read +inf everywhere else.
__global__ void __launch_bounds__(256) gather_wait_done(GatherPeers peers, uint32_t rank, uint32_t world, uint32_t epoch,
                                                        uint32_t T, size_t ld, uint32_t col0, uint32_t col1,
                                                        uint32_t* hostStatus, long long limit) {
    uint32_t* own = peers.flags[rank];
    if (threadIdx.x < world && threadIdx.x != rank)
        spin_until(own + kFlagDone + threadIdx.x, epoch, own + kFlagTimeout, hostStatus, limit);
    __syncthreads();
    if (*reinterpret_cast<volatile uint32_t*>(own + kFlagTimeout) != 0) {
        if (threadIdx.x == 0) {
            *reinterpret_cast<volatile uint32_t*>(hostStatus) = epoch ? epoch : 1u;
            __threadfence_system();
        }
        float*       block = peers.block[rank];
        const float  inf   = __int_as_float(0x7f800000);
        const size_t other = ld - (col1 - col0);
        for (size_t i = threadIdx.x; i < (size_t)T * other; i += blockDim.x) {
            const size_t row = i / other, c = i - row * other;
            block[row * ld + (c < col0 ? c : c + (col1 - col0))] = inf;
        }
    }
    if (threadIdx.x == 0)
        own[kFlagEpoch] = epoch;
}

}  // namespace

int gmm_gather_check(gmm_handle* h, const char* who) {
    if (h->gatherStatus && *reinterpret_cast<volatile uint32_t*>(h->gatherStatus) != 0) {
        char msg[256];
        snprintf(msg, sizeof msg,
                 "%s: the mixture-sharded exchange of call %u on this handle timed out (a peer did not show up); the "
                 "other ranks' columns of that call read +inf.  gmm_gather_reset on every rank clears the condition",
                 who, *h->gatherStatus);
        h->setError(msg);
        return GMM_ERR_TIMEOUT;
    }
    return GMM_OK;
}

/** Collective by contract: every rank calls it, with a host barrier before (no call in flight anywhere) and after. */
int gmm_gather_reset_impl(gmm_handle* h) {
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    if (h->gatherStream)
        GMM_CUDA_TRY(h, cudaStreamSynchronize(h->gatherStream));
    for (auto& b : h->gatherBound) {
        if (b.flags)
            GMM_CUDA_TRY(h, cudaMemsetAsync(b.flags, 0, kFlagWords * sizeof(uint32_t), h->stream));
        b = gmm_handle::GatherBinding();
    }
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    if (h->gatherStatus)
        *reinterpret_cast<volatile uint32_t*>(h->gatherStatus) = 0;
    return GMM_OK;
}

// The join of a started call: the handle's stream waits for the call's last push, then for every peer's DONE.
int gmm_gather_finish_impl(gmm_handle* h, void* d_own_block, size_t flagsOffsetBytes) {
    uint32_t* flags = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(d_own_block) + flagsOffsetBytes);
    for (int i = 0; i < gmm_handle::kGatherBindings; ++i) {
        gmm_handle::GatherBinding& b = h->gatherBound[i];
        if (b.flags != flags || !b.pending)
            continue;
        b.pending = false;
        if (b.world <= 1)
            return GMM_OK;
        GatherPeers peers = {};
        for (uint32_t r = 0; r < b.world; ++r) {
            peers.block[r] = reinterpret_cast<float*>(b.blocks[r]);
            peers.flags[r] = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(b.blocks[r]) + b.flagsOffset);
        }
        const long long spinLimit =
                getenv("GMM_GATHER_TIMEOUT_MS") ? std::max(1ll, atoll(getenv("GMM_GATHER_TIMEOUT_MS"))) * 2000000ll : kSpinLimit;
        GMM_CUDA_TRY(h, cudaStreamWaitEvent(h->stream, h->gatherDone[i], 0));
        gather_wait_done<<<1, 256, 0, h->stream>>>(peers, b.rank, b.world, b.epoch, b.T, b.ld, b.col0, b.col1,
                                                   h->gatherStatusDev, spinLimit);
        h->launches++;
        GMM_CUDA_TRY(h, cudaGetLastError());
        return GMM_OK;
    }
    return GMM_OK;  // nothing pending on this block
}

int gmm_gather_score(gmm_handle* h, uint32_t T, const float* d_x, uint32_t rank, uint32_t world,
                     void* const* d_blocks, size_t ld, uint32_t colOffset, size_t flagsOffsetBytes, uint32_t chunks,
                     bool join) {
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
    if (!h->gatherStatus) {
        GMM_CUDA_TRY(h, cudaHostAlloc(reinterpret_cast<void**>(&h->gatherStatus), 64, cudaHostAllocMapped));
        *h->gatherStatus = 0;
        GMM_CUDA_TRY(h, cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->gatherStatusDev), h->gatherStatus, 0));
    }
    if (int rc = gmm_gather_check(h, "gmm_score_device_gather"))
        return rc;
    // The epoch continues the BLOCK's sequence: a handle that meets a block for the first time reads the block's own
    // epoch word (0 in a fresh block) -- a new engine over used blocks, or a restarted rank, stays in step with its peers.
    gmm_handle::GatherBinding* bound = nullptr;
    for (auto& b : h->gatherBound)
        if (b.flags == peers.flags[rank])
            bound = &b;
    if (!bound) {
        uint32_t words[kFlagWords];
        GMM_CUDA_TRY(h, cudaMemcpyAsync(words, peers.flags[rank], sizeof words, cudaMemcpyDeviceToHost, h->stream));
        GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
        if (words[kFlagTimeout] != 0) {
            h->setError("gmm_score_device_gather: the block carries the marker of a timed-out exchange; gmm_gather_reset "
                        "on every rank (or fresh blocks) first");
            return GMM_ERR_TIMEOUT;
        }
        bound        = &h->gatherBound[h->gatherBoundNext];
        h->gatherBoundNext = (h->gatherBoundNext + 1) % gmm_handle::kGatherBindings;
        if (bound->pending)  // the slot's previous block still has a started call: join it before the slot is reused
            if (int rc = gmm_gather_finish_impl(h, bound->blocks[bound->rank], bound->flagsOffset))
                return rc;
        *bound       = gmm_handle::GatherBinding();
        bound->flags = peers.flags[rank];
        bound->epoch = words[kFlagEpoch];
    }
    if (bound->pending)  // a started call on this block was never joined
        if (int rc = gmm_gather_finish_impl(h, d_blocks[rank], flagsOffsetBytes))
            return rc;
    // bumped only when the call is fully enqueued (below): a call that fails half way is repeated under the same epoch
    const uint32_t epoch = bound->epoch + 1;
    const long long spinLimit =
            getenv("GMM_GATHER_TIMEOUT_MS") ? std::max(1ll, atoll(getenv("GMM_GATHER_TIMEOUT_MS"))) * 2000000ll : kSpinLimit;
    float*         own   = peers.block[rank] + colOffset;
    // The exchange is overlapped range by range.  Short calls are cut into column ranges (whole B tiles over all the
    // frames; only a regular model has its tiles at known columns), long ones into frame ranges (whole 128-row tiles
    // over all the columns): the pushes of a frame range are long contiguous row pieces, which matters with seven
    // peers -- 8192 frames at N=8 in 8 column ranges of 4 KB row pieces took 7.4 ms, in one range 3.7 ms
    // (profiles/sweep_push_n8_T8192_r01.txt) -- while a frame range of fewer than ~1024 frames starves the tensor
    // kernel (the model's B tiles are re-read for every range).
    constexpr uint32_t kTileRows = 128;
    const uint32_t     FT        = (T + kTileRows - 1) / kTileRows;
    static const char* splitEnv  = getenv("GMM_GATHER_SPLIT");
    if (chunks == 0)
        chunks = 1;
    chunks = std::min<uint32_t>(chunks, gmm_handle::kGatherEvents - 1);
    bool byFrames = T / chunks >= 1024;
    if (splitEnv)
        byFrames = !strcmp(splitEnv, "frames");
    if (!byFrames && !tm.regular)
        chunks = 1;
    chunks = std::min<uint32_t>(chunks, byFrames ? FT : std::max(1u, tm.nTiles));
    static const int pushCtas = getenv("GMM_GATHER_CTAS") ? std::max(1, atoi(getenv("GMM_GATHER_CTAS"))) : 128;

    static const int unrollEnv = getenv("GMM_GATHER_UNROLL") ? atoi(getenv("GMM_GATHER_UNROLL")) : 0;
    // measured at N=2 (profiles/sweep_push_n2_*_r01.txt): one load in flight per thread and 128 CTAs disturb the tensor
    // kernel running beside the push least (T=8192, 8 ranges: 0.951 ms; four loads in flight: 1.065 ms)
    const int        unroll    = unrollEnv ? unrollEnv : 1;

    // Frame ranges: range c of n takes n - c parts of n(n+1)/2 -- what the call pays for the exchange is the push of
    // the LAST range, so that one is kept small (N=2, T=8192, 4 ranges: 0.995 -> 0.947 ms).  Column ranges stay even:
    // the same cut made short calls slower (T=1024, 2 ranges: 0.168 -> 0.180 ms; profiles/sweep_push_n2_*_uneven_r01.txt).
    static const bool evenEnv    = getenv("GMM_GATHER_EVEN") != nullptr;
    const bool        evenRanges = evenEnv || !byFrames;
    auto cut = [&](uint32_t total, uint32_t c) -> uint32_t {  // first unit of range c; every range keeps >= 1 unit
        if (c >= chunks)
            return total;
        uint64_t b;
        if (evenRanges)
            b = (uint64_t)total * c / chunks;
        else {
            const uint64_t done = (uint64_t)c * chunks - (uint64_t)c * (c - 1) / 2;  // sum of (n - i), i < c
            b                   = (uint64_t)total * done / ((uint64_t)chunks * (chunks + 1) / 2);
        }
        b = std::max<uint64_t>(b, c);
        return (uint32_t)std::min<uint64_t>(b, total - (chunks - c));
    };

    bool readyWaited = false;  // copy-engine pushes: the READY wait kernel runs once per call
    gather_signal_ready<<<1, 32, 0, h->stream>>>(peers, rank, world, epoch);
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    for (uint32_t c = 0; c < chunks; ++c) {
        uint32_t row0 = 0, rows = T, col0 = colOffset, col1 = colOffset + h->M;
        int      rc;
        if (byFrames) {
            row0 = cut(FT, c) * kTileRows;
            rows = std::min<uint32_t>(T, cut(FT, c + 1) * kTileRows) - row0;
            // all frames are prepared by the first range's call: with a preparation kernel between the ranges the
            // push CTAs settle on the SMs first and the next range's scoring CTAs (all of an SM's shared memory)
            // wait for them -- measured: no overlap at all
            rc = gmm_tc_score_tiles(h, rows, c == 0 ? d_x : d_x + (size_t)row0 * h->D, own + (size_t)row0 * ld, ld, 0,
                                    tm.nTiles, c == 0, c == 0 ? T : 0, row0 / kTileRows);
        }
        else {
            const uint32_t t0 = cut(tm.nTiles, c), t1 = cut(tm.nTiles, c + 1);
            rc = gmm_tc_score_tiles(h, T, d_x, own, ld, t0, t1, c == 0);
            if (chunks > 1) {
                col0 = colOffset + t0 * tm.mixPerTile;
                col1 = t1 == tm.nTiles ? colOffset + h->M : colOffset + t1 * tm.mixPerTile;
            }
        }
        if (rc)
            return rc;
        if (world > 1) {
            GMM_CUDA_TRY(h, cudaEventRecord(h->gatherEvent[c], h->stream));
            GMM_CUDA_TRY(h, cudaStreamWaitEvent(h->gatherStream, h->gatherEvent[c], 0));
            bool vec = ld % 4 == 0 && col0 % 4 == 0 && (col1 - col0) % 4 == 0;
            for (uint32_t r = 0; r < world && vec; ++r)
                vec = reinterpret_cast<uintptr_t>(peers.block[r]) % 16 == 0;
            const int last = c + 1 == chunks;
            // With ONE peer, ranges of at least 4 MB go through the copy engine (GMM_GATHER_COPY=sm|ce forces either).  With
            // more peers the copies of a range measured slower than the push kernel, one after the other on this stream
            // (N = 4, 8192 frames: 4.03 against 3.04-3.46 ms) as well as side by side on a stream per peer (3.96-4.55 ms;
            // profiles/bench_n4_ce_{serial,parallel}_r02.json): the copy engines deliver ~550 GB/s in total, the push
            // kernel 650 once the exchange and not the scoring is what the call waits for.
            const char* copyEnv = getenv("GMM_GATHER_COPY");
            const size_t rangeBytes = (size_t)rows * (col1 - col0) * 4;
            const bool   useCE = copyEnv ? !strcmp(copyEnv, "ce") : (world == 2 && rangeBytes >= (4u << 20));
            if (useCE) {
                if (!readyWaited) {
                    gather_wait_ready<<<1, 32, 0, h->gatherStream>>>(peers, rank, world, epoch, h->gatherStatusDev, spinLimit);
                    h->launches++;
                    readyWaited = true;
                }
                const size_t off = (size_t)row0 * ld + col0;
                for (uint32_t k = 1; k < world; ++k) {
                    const uint32_t r = (rank + k) % world;
                    GMM_CUDA_TRY(h, cudaMemcpy2DAsync(peers.block[r] + off, ld * 4, peers.block[rank] + off, ld * 4,
                                                      (size_t)(col1 - col0) * 4, rows, cudaMemcpyDeviceToDevice,
                                                      h->gatherStream));
                }
                if (last) {
                    gather_announce_done<<<1, 32, 0, h->gatherStream>>>(peers, rank, world, epoch);
                    h->launches++;
                }
                GMM_CUDA_TRY(h, cudaGetLastError());
                continue;
            }
#define GMM_PUSH(V, U)                                                                                             \
    gather_push_kernel<V, U><<<pushCtas, 256, 0, h->gatherStream>>>(peers, rank, world, row0, rows, ld, col0, col1, \
                                                                    epoch, last, h->gatherStatusDev, spinLimit)
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
    const int slot = (int)(bound - h->gatherBound);
    if (world > 1) {
        if (!h->gatherDone[slot])
            GMM_CUDA_TRY(h, cudaEventCreateWithFlags(&h->gatherDone[slot], cudaEventDisableTiming));
        GMM_CUDA_TRY(h, cudaEventRecord(h->gatherDone[slot], h->gatherStream));
    }
    bound->epoch   = epoch;
    bound->pending = true;
    bound->rank = rank, bound->world = world, bound->T = T, bound->ld = ld, bound->flagsOffset = flagsOffsetBytes;
    bound->col0 = colOffset, bound->col1 = colOffset + h->M;
    for (uint32_t r = 0; r < 8; ++r)
        bound->blocks[r] = r < world ? d_blocks[r] : nullptr;
    // begin only: the caller joins with gmm_gather_finish -- after it has started the NEXT call on another block, so that
    // this call's last push runs under the next call's scoring
    return join ? gmm_gather_finish_impl(h, d_blocks[rank], flagsOffsetBytes) : GMM_OK;
}
