// C-ABI of libgmm_b200.so (include/gmm_b200.h): handle lifetime, model upload / packing, and the host- and
// device-pointer scoring entry points.  No CPU fallback anywhere: every scoring call ends in a CUDA kernel.
#include <math.h>
#include <sched.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>

#include "gmm_internal.cuh"

static std::string g_createError;

// GMM_TRACE: device-side timeline of the pipelined host entry points (events on the lanes' streams, printed per call)
struct TraceMark {
    cudaEvent_t ev;
    int         group, lane;
    const char* what;
};
static std::vector<TraceMark> g_trace;
static cudaEvent_t            g_traceStart = nullptr;
static void traceMark(gmm_handle* h, int group, const char* what) {
    static const bool on = getenv("GMM_TRACE") != nullptr;
    if (!on || !g_traceStart)
        return;
    TraceMark m;
    cudaEventCreate(&m.ev);
    cudaEventRecord(m.ev, h->stream);
    m.group = group, m.lane = h->lane, m.what = what;
    g_trace.push_back(m);
}

int gmm_handle::devAlloc(void** p, size_t bytes, bool quant) {
    *p = nullptr;
    if (bytes == 0)
        bytes = 16;
    GMM_CUDA_TRY(this, cudaMalloc(p, bytes));
    (quant ? ownedQuant : owned).push_back(*p);
    return GMM_OK;
}

int gmm_handle::upload(void** p, const void* src, size_t bytes, bool quant) {
    int rc = devAlloc(p, bytes, quant);
    if (rc)
        return rc;
    if (bytes)
        GMM_CUDA_TRY(this, cudaMemcpyAsync(*p, src, bytes, cudaMemcpyHostToDevice, stream));
    return GMM_OK;
}

int gmm_handle::reserve(int slot, size_t bytes, void** out) {
    if (bytes > scratchBytes[slot]) {
        GMM_CUDA_TRY(this, cudaStreamSynchronize(stream));
        if (scratch[slot])
            GMM_CUDA_TRY(this, cudaFree(scratch[slot]));
        scratch[slot]      = nullptr;
        scratchBytes[slot] = 0;
        size_t want        = bytes + bytes / 4 + 256;
        GMM_CUDA_TRY(this, cudaMalloc(&scratch[slot], want));
        scratchBytes[slot] = want;
    }
    *out = scratch[slot];
    return GMM_OK;
}

HostPool& gmm_handle::hostPool() {
    if (!pool) {
        int         n   = 4;
        const char* env = getenv("GMM_HOST_THREADS");
        cpu_set_t   set;
        if (env)
            n = std::max(1, atoi(env));
        else if (sched_getaffinity(0, sizeof(set), &set) == 0)
            n = std::max(1, std::min(4, CPU_COUNT(&set) / 2));
        pool = new HostPool(n - 1);
        workerWords.resize(n);
        workerStamps.resize(n);
    }
    return *pool;
}

static __global__ void __launch_bounds__(256) table_copy_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src,
                                                                size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = src[i];
}

int gmm_copy_tables(gmm_handle* h, void* d_dst, const void* pinnedSrc, size_t bytes) {
    static const bool dma = [] {
        const char* e = getenv("GMM_TABLE_COPY");
        return e && !strcmp(e, "dma");
    }();
    if (bytes == 0)
        return GMM_OK;
    if (dma || (bytes & 15) || ((uintptr_t)d_dst & 15) || ((uintptr_t)pinnedSrc & 15)) {
        GMM_CUDA_TRY(h, cudaMemcpyAsync(d_dst, pinnedSrc, bytes, cudaMemcpyHostToDevice, h->stream));
        return GMM_OK;
    }
    const size_t n    = bytes / 16;
    const int    grid = (int)std::min<size_t>((n + 255) / 256, (size_t)h->smCount * 4);
    table_copy_kernel<<<grid, 256, 0, h->stream>>>((uint4*)d_dst, (const uint4*)pinnedSrc, n);
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    return GMM_OK;
}

int gmm_handle::pinnedCopied(int slot) {
    if (!pinEvent[slot])
        GMM_CUDA_TRY(this, cudaEventCreateWithFlags(&pinEvent[slot], cudaEventDisableTiming));
    GMM_CUDA_TRY(this, cudaEventRecord(pinEvent[slot], stream));
    return GMM_OK;
}

int gmm_handle::reservePinned(int slot, size_t bytes, void** out) {
    if (pinEvent[slot])
        GMM_CUDA_TRY(this, cudaEventSynchronize(pinEvent[slot]));  // the previous copy out of this block is done
    if (bytes > pinnedBytes[slot]) {
        GMM_CUDA_TRY(this, cudaStreamSynchronize(stream));
        if (pinned[slot])
            GMM_CUDA_TRY(this, cudaFreeHost(pinned[slot]));
        pinned[slot]      = nullptr;
        pinnedBytes[slot] = 0;
        const size_t want = bytes + bytes / 4 + 4096;
        GMM_CUDA_TRY(this, cudaHostAlloc(&pinned[slot], want, cudaHostAllocDefault));
        pinnedBytes[slot] = want;
    }
    *out = pinned[slot];
    return GMM_OK;
}

int gmm_handle::useLane(int which) {
    if (which == lane)
        return GMM_OK;
    if (which < 0 || which >= kMaxLanes)
        return GMM_ERR_ARG;
    if (!laneEvent)
        GMM_CUDA_TRY(this, cudaEventCreateWithFlags(&laneEvent, cudaEventDisableTiming));
    LaneState& out = parked[lane];  // park the active lane ...
    LaneState& in  = parked[which];  // ... and take the other one's state
    if (which != 0 && !in.stream)
        GMM_CUDA_TRY(this, cudaStreamCreateWithFlags(&in.stream, cudaStreamNonBlocking));
    out.stream = stream;
    stream     = in.stream;
    in.stream  = nullptr;
    for (int i = 0; i < SCR_COUNT; ++i) {
        out.scratch[i] = scratch[i], out.scratchBytes[i] = scratchBytes[i];
        scratch[i] = in.scratch[i], scratchBytes[i] = in.scratchBytes[i];
        in.scratch[i] = nullptr, in.scratchBytes[i] = 0;
    }
    for (int i = 0; i < 2; ++i) {
        out.pinned[i] = pinned[i], out.pinnedBytes[i] = pinnedBytes[i], out.pinEvent[i] = pinEvent[i];
        pinned[i] = in.pinned[i], pinnedBytes[i] = in.pinnedBytes[i], pinEvent[i] = in.pinEvent[i];
        in.pinned[i] = nullptr, in.pinnedBytes[i] = 0, in.pinEvent[i] = nullptr;
    }
    lane = which;
    return GMM_OK;
}

static void releaseFloatModel(gmm_handle* h) {
    cudaStreamSynchronize(h->stream);
    for (void* p : h->owned)
        cudaFree(p);
    h->owned.clear();
    h->floatValid = false;
    h->direct     = DirectModel();
    h->tensor     = TensorModel();
    h->accum      = AccumModel();  // its arrays were in `owned`
}

static void releaseQuantModel(gmm_handle* h) {
    cudaStreamSynchronize(h->stream);
    for (void* p : h->ownedQuant)
        cudaFree(p);
    h->ownedQuant.clear();
    h->quant   = QuantModel();
    h->qtensor = QTensorModel();
}

// Host frames -> device x [T x D] in SCR_X.  Plain copy, or, with a front end set (gmm_set_lda), the narrow base
// features [T x baseDim] go up and the device builds the windows and applies the matrix (gmm_lda.cu).
// frameOffsets: nSeg + 1 entries starting at 0 (windows do not cross segment boundaries).
static int uploadFrames(gmm_handle* h, uint32_t D, uint32_t nSeg, const uint32_t* frameOffsets, const float* x, void** dx) {
    const uint32_t T = frameOffsets[nSeg];
    int            rc;
    if ((rc = h->reserve(SCR_X, (size_t)std::max(1u, T) * D * 4, dx)))
        return rc;
    if (T == 0)
        return GMM_OK;
    if (!h->lda.valid) {
        GMM_CUDA_TRY(h, cudaMemcpyAsync(*dx, x, (size_t)T * D * 4, cudaMemcpyHostToDevice, h->stream));
        return GMM_OK;
    }
    if (h->lda.D != D) {
        h->setError("front end (gmm_set_lda) produces a different dimension than the model takes");
        return GMM_ERR_ARG;
    }
    void* db = nullptr;
    if ((rc = h->reserve(SCR_BASE, (size_t)T * h->lda.baseDim * 4, &db)))
        return rc;
    GMM_CUDA_TRY(h, cudaMemcpyAsync(db, x, (size_t)T * h->lda.baseDim * 4, cudaMemcpyHostToDevice, h->stream));
    return gmm_lda_launch(h, nSeg, frameOffsets, (const float*)db, (float*)*dx);
}

extern "C" {

const char* gmm_version(void) {
    return "gmm_b200 0.1 (sm_100a)";
}

int gmm_create(gmm_handle** out, int device) {
    if (!out) {
        g_createError = "gmm_create: out is NULL";
        return GMM_ERR_ARG;
    }
    *out  = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_createError = std::string("gmm_create: no CUDA device (") + cudaGetErrorString(e) + ")";
        return GMM_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= n) {
        g_createError = "gmm_create: device index out of range";
        return GMM_ERR_ARG;
    }
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) {
        g_createError = std::string("gmm_create: ") + cudaGetErrorString(e);
        return GMM_ERR_CUDA;
    }
    if (prop.major != 10) {
        g_createError = "gmm_create: device is not sm_100 (the library is built for sm_100a only)";
        return GMM_ERR_NO_DEVICE;
    }
    e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        g_createError = std::string("gmm_create: ") + cudaGetErrorString(e);
        return GMM_ERR_CUDA;
    }
    gmm_handle* h = new gmm_handle;
    h->device     = device;
    h->smCount    = prop.multiProcessorCount;
    e             = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        g_createError = std::string("gmm_create: ") + cudaGetErrorString(e);
        delete h;
        return GMM_ERR_CUDA;
    }
    h->ownStream = true;
    *out         = h;
    return GMM_OK;
}

void gmm_destroy(gmm_handle* h) {
    if (!h)
        return;
    cudaSetDevice(h->device);
    releaseFloatModel(h);
    releaseQuantModel(h);
    for (size_t i = 0; i < h->timed.size(); ++i) {
        cudaEventDestroy(h->timed[i].first);
        cudaEventDestroy(h->timed[i].second);
    }
    h->useLane(0);
    for (int i = 0; i < SCR_COUNT; ++i)
        if (h->scratch[i])
            cudaFree(h->scratch[i]);
    for (int i = 0; i < 2; ++i) {
        if (h->pinned[i])
            cudaFreeHost(h->pinned[i]);
        if (h->pinEvent[i])
            cudaEventDestroy(h->pinEvent[i]);
    }
    for (int l = 1; l < gmm_handle::kMaxLanes; ++l) {
        gmm_handle::LaneState& ls = h->parked[l];
        for (int i = 0; i < SCR_COUNT; ++i)
            if (ls.scratch[i])
                cudaFree(ls.scratch[i]);
        for (int i = 0; i < 2; ++i) {
            if (ls.pinned[i])
                cudaFreeHost(ls.pinned[i]);
            if (ls.pinEvent[i])
                cudaEventDestroy(ls.pinEvent[i]);
        }
        if (ls.stream)
            cudaStreamDestroy(ls.stream);
    }
    delete h->pool;
    if (h->gatherStream) {
        cudaStreamDestroy(h->gatherStream);
        for (int i = 0; i < gmm_handle::kGatherEvents; ++i)
            if (h->gatherEvent[i])
                cudaEventDestroy(h->gatherEvent[i]);
    }
    for (cudaEvent_t e : h->gatherDone)
        if (e)
            cudaEventDestroy(e);
    if (h->gatherStatus)
        cudaFreeHost(h->gatherStatus);
    if (h->laneEvent)
        cudaEventDestroy(h->laneEvent);
    if (h->copyStream)
        cudaStreamDestroy(h->copyStream);
    for (cudaEvent_t e : h->copyEvent)
        if (e)
            cudaEventDestroy(e);
    for (cudaEvent_t e : h->asyncEvent)
        if (e)
            cudaEventDestroy(e);
    if (h->lda.matrix)
        cudaFree(h->lda.matrix);
    if (h->ownStream && h->stream)
        cudaStreamDestroy(h->stream);
    delete h;
}

const char* gmm_last_error(const gmm_handle* h) {
    return h ? h->error.c_str() : g_createError.c_str();
}

int gmm_set_stream(gmm_handle* h, void* cuda_stream) {
    if (!h)
        return GMM_ERR_ARG;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    if (h->ownStream && h->stream)
        cudaStreamDestroy(h->stream);
    if (cuda_stream) {
        h->stream    = (cudaStream_t)cuda_stream;
        h->ownStream = false;
    }
    else {
        GMM_CUDA_TRY(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        h->ownStream = true;
    }
    return GMM_OK;
}

int gmm_sync(gmm_handle* h) {
    if (!h)
        return GMM_ERR_ARG;
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    for (int i = 0; i < gmm_handle::kAsyncLanes; ++i)  // outstanding gmm_score_async calls
        if (h->asyncToken[i])
            GMM_CUDA_TRY(h, cudaEventSynchronize(h->asyncEvent[i]));
    return GMM_OK;
}

int gmm_quant_kernel(const gmm_handle* h) {
    if (!h || !h->quant.valid)
        return 0;
    return h->qtensor.valid ? 2 : 1;
}

int gmm_kernel_timing(gmm_handle* h, int on) {
    if (!h)
        return GMM_ERR_ARG;
    h->timing = on != 0;
    return GMM_OK;
}

int gmm_kernel_time(gmm_handle* h, double* total_ms, uint64_t* launches) {
    if (!h)
        return GMM_ERR_ARG;
    GMM_CUDA_TRY(h, cudaSetDevice(h->device));
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    double sum = 0.0;
    for (size_t i = 0; i < h->timed.size(); ++i) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, h->timed[i].first, h->timed[i].second) == cudaSuccess)
            sum += ms;
        cudaEventDestroy(h->timed[i].first);
        cudaEventDestroy(h->timed[i].second);
    }
    if (total_ms)
        *total_ms = sum;
    if (launches)
        *launches = h->timed.size();
    h->timed.clear();
    return GMM_OK;
}

uint64_t gmm_launch_count(const gmm_handle* h) {
    return h ? h->launches : 0;
}

int gmm_float_kernel(const gmm_handle* h) {
    return (h && h->floatValid) ? h->kernel : 0;
}

int gmm_float_split(const gmm_handle* h) {
    if (!h || !h->floatValid || h->kernel != GMM_KERNEL_TENSOR || !h->tensor.valid)
        return 0;
    return h->tensor.fp8 ? 2 : 3;
}

int gmm_quant_final_score_rule(gmm_handle* h, int rule) {
    if (!h || (rule != GMM_QUANT_RULE_SIMD && rule != GMM_QUANT_RULE_BATCH_INT)) {
        if (h)
            h->setError("gmm_quant_final_score_rule: unknown rule");
        return GMM_ERR_ARG;
    }
    if (!h->quant.valid) {
        h->setError("gmm_quant_final_score_rule: no quantized model uploaded");
        return GMM_ERR_ARG;
    }
    h->quant.scoreRule = rule;
    return GMM_OK;
}

int gmm_float_accuracy_ok(const gmm_handle* h) {
    if (!h || !h->floatValid)
        return 0;
    return h->kernel != GMM_KERNEL_TENSOR || h->tensor.accuracyOk;
}

void* gmm_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 16) != cudaSuccess)
        return nullptr;
    return p;
}

void gmm_host_free(void* p) {
    if (p)
        cudaFreeHost(p);
}

static int checkOffsets(gmm_handle* h, uint32_t M, uint32_t G, const uint32_t* offsets) {
    if (!offsets || offsets[0] != 0 || offsets[M] != G) {
        h->setError("offsets must be a prefix sum with offsets[0] == 0 and offsets[M] == G");
        return GMM_ERR_ARG;
    }
    for (uint32_t m = 0; m < M; ++m)
        if (offsets[m + 1] < offsets[m]) {
            h->setError("offsets must be non-decreasing");
            return GMM_ERR_ARG;
        }
    return GMM_OK;
}

int gmm_upload_float(gmm_handle* h, uint32_t M, uint32_t G, uint32_t D, uint32_t C, const uint32_t* offsets,
                     const float* mean, const float* isr, const uint32_t* cov_idx, const float* w2,
                     const float* log_norm, int mode, int kernel) {
    if (!h)
        return GMM_ERR_ARG;
    cudaSetDevice(h->device);
    if (D == 0 || C == 0 || (G && (!mean || !cov_idx || !w2)) || !isr || !log_norm ||
        (mode != GMM_MODE_MAX && mode != GMM_MODE_SUM) || kernel < GMM_KERNEL_AUTO || kernel > GMM_KERNEL_TENSOR) {
        h->setError("gmm_upload_float: invalid argument");
        return GMM_ERR_ARG;
    }
    int rc = checkOffsets(h, M, G, offsets);
    if (rc)
        return rc;
    for (uint32_t g = 0; g < G; ++g)
        if (cov_idx[g] >= C) {
            h->setError("gmm_upload_float: covariance index out of range");
            return GMM_ERR_ARG;
        }
    releaseFloatModel(h);  // a handle holds one float model; re-upload replaces it
    h->M    = M;
    h->G    = G;
    h->D    = D;
    h->C    = C;
    h->mode = mode;
    h->maxK = 0;
    for (uint32_t m = 0; m < M; ++m)
        h->maxK = std::max(h->maxK, offsets[m + 1] - offsets[m]);

    // direct layout: rows padded to a multiple of 4 floats (16/32/48/64/96/128 kernel variants)
    uint32_t DP = D <= 16 ? 16 : D <= 32 ? 32 : D <= 48 ? 48 : D <= 64 ? 64 : D <= 96 ? 96 : 128;
    if (D > 128) {
        h->setError("feature dimension above 128 is not supported");
        return GMM_ERR_UNSUPPORTED;
    }
    DirectModel& dm = h->direct;
    dm.DP           = DP;
    {
        std::vector<float> pm((size_t)G * DP, 0.f), pi((size_t)C * DP, 0.f);
        for (size_t g = 0; g < G; ++g)
            memcpy(&pm[g * DP], mean + g * D, D * sizeof(float));
        for (size_t c = 0; c < C; ++c)
            memcpy(&pi[c * DP], isr + c * D, D * sizeof(float));
        if ((rc = h->upload((void**)&dm.mean, pm.data(), pm.size() * sizeof(float))))
            return rc;
        if ((rc = h->upload((void**)&dm.isr, pi.data(), pi.size() * sizeof(float))))
            return rc;
        GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));  // staging vectors go out of scope
    }
    if ((rc = h->upload((void**)&dm.covIdx, cov_idx, (size_t)G * 4)))
        return rc;
    if ((rc = h->upload((void**)&dm.w2, w2, (size_t)G * 4)))
        return rc;
    if ((rc = h->upload((void**)&dm.logNorm, log_norm, (size_t)C * 4)))
        return rc;
    if ((rc = h->upload((void**)&dm.offsets, offsets, (size_t)(M + 1) * 4)))
        return rc;
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));

    h->kernel = GMM_KERNEL_DIRECT;
    if (kernel != GMM_KERNEL_DIRECT) {
        rc = gmm_tc_pack(h, offsets, mean, isr, cov_idx, w2, log_norm, kernel == GMM_KERNEL_TENSOR);
        if (rc == GMM_OK && h->tensor.valid)
            h->kernel = GMM_KERNEL_TENSOR;
        else if (kernel == GMM_KERNEL_TENSOR) {
            if (rc == GMM_OK) {
                h->setError("gmm_upload_float: model does not fit the tensor kernel");
                rc = GMM_ERR_UNSUPPORTED;
            }
            return rc;
        }
        else if (rc != GMM_OK && rc != GMM_ERR_UNSUPPORTED)
            return rc;
    }
    h->floatValid = true;
    return GMM_OK;
}

int gmm_score_device(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->floatValid) {
        h->setError("gmm_score: no float model uploaded");
        return GMM_ERR_ARG;
    }
    if (ld < h->M || (T && (!d_x || !d_scores))) {
        h->setError("gmm_score: invalid argument (ld < M or NULL buffer)");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    if (h->kernel == GMM_KERNEL_TENSOR)
        return gmm_tc_score(h, T, d_x, d_scores, d_best, ld);
    return gmm_direct_score(h, T, d_x, d_scores, d_best, ld);
}

// argument checks of the mixture-sharded calls: < 0 = error, 1 = nothing to do (no frames), 0 = go
static int gatherArgs(gmm_handle* h, uint32_t T, const float* d_x, uint32_t rank, uint32_t world, void* const* d_blocks,
                      size_t ld, uint32_t colOffset, size_t flagsOffsetBytes) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->floatValid || h->kernel != GMM_KERNEL_TENSOR || !h->tensor.valid) {
        h->setError("gmm_score_device_gather: needs a float model on the tensor kernel");
        return GMM_ERR_UNSUPPORTED;
    }
    if (world == 0 || world > 8 || rank >= world || !d_blocks || flagsOffsetBytes % 128 || (size_t)colOffset + h->M > ld ||
        flagsOffsetBytes < (size_t)T * ld * 4) {
        h->setError("gmm_score_device_gather: 1..8 ranks, rank < world, the shard's columns inside ld, flag words "
                    "128-byte aligned behind the score rows");
        return GMM_ERR_ARG;
    }
    if (T > 32768) {
        h->setError("gmm_score_device_gather: at most 32768 frames per call (one launch)");
        return GMM_ERR_ARG;
    }
    if (T == 0)
        return 1;
    if (!d_x) {
        h->setError("gmm_score_device_gather: NULL frames");
        return GMM_ERR_ARG;
    }
    return 0;
}

int gmm_score_device_gather(gmm_handle* h, uint32_t T, const float* d_x, uint32_t rank, uint32_t world,
                            void* const* d_blocks, size_t ld, uint32_t colOffset, size_t flagsOffsetBytes,
                            uint32_t chunks) {
    if (int rc = gatherArgs(h, T, d_x, rank, world, d_blocks, ld, colOffset, flagsOffsetBytes))
        return rc < 0 ? rc : GMM_OK;
    cudaSetDevice(h->device);
    return gmm_gather_score(h, T, d_x, rank, world, d_blocks, ld, colOffset, flagsOffsetBytes, chunks);
}

int gmm_score_device_gather_begin(gmm_handle* h, uint32_t T, const float* d_x, uint32_t rank, uint32_t world,
                                  void* const* d_blocks, size_t ld, uint32_t colOffset, size_t flagsOffsetBytes,
                                  uint32_t chunks) {
    if (int rc = gatherArgs(h, T, d_x, rank, world, d_blocks, ld, colOffset, flagsOffsetBytes))
        return rc < 0 ? rc : GMM_OK;
    cudaSetDevice(h->device);
    return gmm_gather_score(h, T, d_x, rank, world, d_blocks, ld, colOffset, flagsOffsetBytes, chunks, false);
}

int gmm_gather_finish(gmm_handle* h, void* d_own_block, size_t flagsOffsetBytes) {
    if (!h || !d_own_block)
        return GMM_ERR_ARG;
    cudaSetDevice(h->device);
    return gmm_gather_finish_impl(h, d_own_block, flagsOffsetBytes);
}

int gmm_gather_status(gmm_handle* h, uint32_t* timedOutCall) {
    if (!h || !timedOutCall)
        return GMM_ERR_ARG;
    *timedOutCall = h->gatherStatus ? *reinterpret_cast<volatile uint32_t*>(h->gatherStatus) : 0;
    return GMM_OK;
}

int gmm_gather_reset(gmm_handle* h) {
    if (!h)
        return GMM_ERR_ARG;
    cudaSetDevice(h->device);
    return gmm_gather_reset_impl(h);
}

// a block that is allocated, mapped or released may sit where another one was: forget what was read from blocks
static void forgetGatherBlocks(gmm_handle* h) {
    for (auto& b : h->gatherBound)
        b = gmm_handle::GatherBinding();
}

int gmm_peer_alloc(gmm_handle* h, size_t bytes, void** d_ptr, void* ipcHandle) {
    if (!h || !d_ptr || !ipcHandle)
        return GMM_ERR_ARG;
    forgetGatherBlocks(h);
    cudaSetDevice(h->device);
    *d_ptr = nullptr;
    GMM_CUDA_TRY(h, cudaMalloc(d_ptr, bytes ? bytes : 16));
    GMM_CUDA_TRY(h, cudaMemset(*d_ptr, 0, bytes ? bytes : 16));  // the flag words of a gathered block start at 0
    cudaIpcMemHandle_t hd;
    GMM_CUDA_TRY(h, cudaIpcGetMemHandle(&hd, *d_ptr));
    memcpy(ipcHandle, &hd, sizeof(hd));
    return GMM_OK;
}

int gmm_peer_open(gmm_handle* h, const void* ipcHandle, void** d_ptr) {
    if (!h || !d_ptr || !ipcHandle)
        return GMM_ERR_ARG;
    forgetGatherBlocks(h);
    cudaSetDevice(h->device);
    cudaIpcMemHandle_t hd;
    memcpy(&hd, ipcHandle, sizeof(hd));
    GMM_CUDA_TRY(h, cudaIpcOpenMemHandle(d_ptr, hd, cudaIpcMemLazyEnablePeerAccess));
    return GMM_OK;
}

int gmm_peer_close(gmm_handle* h, void* d_ptr) {
    if (!h)
        return GMM_ERR_ARG;
    forgetGatherBlocks(h);
    cudaSetDevice(h->device);
    GMM_CUDA_TRY(h, cudaIpcCloseMemHandle(d_ptr));
    return GMM_OK;
}

int gmm_peer_free(gmm_handle* h, void* d_ptr) {
    if (!h)
        return GMM_ERR_ARG;
    forgetGatherBlocks(h);
    cudaSetDevice(h->device);
    GMM_CUDA_TRY(h, cudaFree(d_ptr));
    return GMM_OK;
}

int gmm_score_subset_device(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const uint32_t* listOffsets,
                            const uint32_t* mixtures, const float* d_x, float* d_scores, size_t ld) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->floatValid) {
        h->setError("gmm_score_subset: no float model uploaded");
        return GMM_ERR_ARG;
    }
    if (nSeg == 0)
        return GMM_OK;
    if (!frameOffsets || !listOffsets || (listOffsets[nSeg] && !mixtures) || (frameOffsets[nSeg] && (!d_x || !d_scores))) {
        h->setError("gmm_score_subset: NULL argument");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    if (!gmm_tc_subset_capable(h)) {
        h->setError("gmm_score_subset: needs the tensor kernel and a model whose mixtures all have the same number "
                    "of (live) densities; score all mixtures instead");
        return GMM_ERR_UNSUPPORTED;
    }
    return gmm_tc_score_subset(h, nSeg, frameOffsets, listOffsets, mixtures, d_x, d_scores, ld);
}

uint32_t gmm_score_subset_ld(const gmm_handle* h, uint32_t maxListLength) {
    return (h && h->floatValid && h->tensor.valid) ? gmm_tc_subset_ld(h, maxListLength) : 0;
}

static int copyOut(gmm_handle* h, void* host, const void* dev, uint32_t T, uint32_t M, size_t ld, size_t devLd,
                   cudaStream_t stream = nullptr) {
    if (!stream)
        stream = h->stream;
    if (ld == devLd)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(host, dev, (size_t)T * ld * 4, cudaMemcpyDeviceToHost, stream));
    else
        GMM_CUDA_TRY(h, cudaMemcpy2DAsync(host, ld * 4, dev, devLd * 4, (size_t)M * 4, T, cudaMemcpyDeviceToHost, stream));
    return GMM_OK;
}

// Host frames in, host scores out, everything enqueued and nothing waited for: the body of gmm_score_host and
// gmm_score_async.  The frames go up in one copy; the batch is scored in up to kCopyEvents chunks on the handle's
// stream, and every finished chunk's rows go back on a copy stream of their own while the next chunk is scored (the
// device-to-host copy of 4 * ld bytes per frame is what bounds the call: PCIe).  On return `done` (an
// event on the copy stream, or the handle's stream for a single chunk) marks the last copy.
static int scoreHostEnqueue(gmm_handle* h, uint32_t T, const float* x, float* scores, uint32_t* best, size_t ld,
                            cudaStream_t* doneStream) {
    void *dx = nullptr, *ds = nullptr, *db = nullptr;
    int   rc;
    // device rows: the caller's pitch when it keeps row pieces 16-byte aligned (then one contiguous copy per chunk),
    // else M rounded up to a multiple of 4 -- the tensor kernel stores aligned row pieces straight from registers
    const size_t   devLd  = (ld % 4 == 0 && ld <= (size_t)h->M + 1024) ? ld : ((size_t)h->M + 3) / 4 * 4;
    const uint32_t one[2] = {0, T};
    if ((rc = uploadFrames(h, h->D, 1, one, x, &dx)))
        return rc;
    if ((rc = h->reserve(SCR_SCORE, (size_t)T * devLd * 4, &ds)))
        return rc;
    if (best && (rc = h->reserve(SCR_BEST, (size_t)T * devLd * 4, &db)))
        return rc;
    uint32_t nChunks = 1;
    {
        const char* e = getenv("GMM_HOST_CHUNKS");
        nChunks       = e ? (uint32_t)std::max(1, atoi(e)) : (T >= 8192 ? std::min<uint32_t>(8, T / 4096) : 1);
        nChunks       = std::min<uint32_t>(nChunks, gmm_handle::kCopyEvents);
    }
    if (nChunks > 1 && !h->copyStream)
        GMM_CUDA_TRY(h, cudaStreamCreateWithFlags(&h->copyStream, cudaStreamNonBlocking));
    const uint32_t per = ((T + nChunks - 1) / nChunks + 127) / 128 * 128;
    uint32_t       c   = 0;
    for (uint32_t t0 = 0; t0 < T; t0 += per, ++c) {
        const uint32_t n = std::min(per, T - t0);
        const size_t   o = (size_t)t0 * devLd;
        if ((rc = gmm_score_device(h, n, (const float*)dx + (size_t)t0 * h->D, (float*)ds + o,
                                   db ? (uint32_t*)db + o : nullptr, devLd)))
            return rc;
        cudaStream_t cs = h->stream;
        if (nChunks > 1) {
            if (!h->copyEvent[c])
                GMM_CUDA_TRY(h, cudaEventCreateWithFlags(&h->copyEvent[c], cudaEventDisableTiming));
            GMM_CUDA_TRY(h, cudaEventRecord(h->copyEvent[c], h->stream));
            GMM_CUDA_TRY(h, cudaStreamWaitEvent(h->copyStream, h->copyEvent[c], 0));
            cs = h->copyStream;
        }
        if ((rc = copyOut(h, scores + (size_t)t0 * ld, (float*)ds + o, n, h->M, ld, devLd, cs)))
            return rc;
        if (best && (rc = copyOut(h, best + (size_t)t0 * ld, (uint32_t*)db + o, n, h->M, ld, devLd, cs)))
            return rc;
    }
    if (nChunks > 1) {
        // later work on the handle's stream (it may overwrite the scratch) comes after the copies
        cudaEvent_t& e = h->copyEvent[gmm_handle::kCopyEvents];
        if (!e)
            GMM_CUDA_TRY(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        GMM_CUDA_TRY(h, cudaEventRecord(e, h->copyStream));
        GMM_CUDA_TRY(h, cudaStreamWaitEvent(h->stream, e, 0));
    }
    *doneStream = h->stream;
    return GMM_OK;
}

static int checkScoreHostArgs(gmm_handle* h, uint32_t T, const float* x, float* scores, size_t ld, const char* who) {
    if (!h->floatValid) {
        h->setError(std::string(who) + ": no float model uploaded");
        return GMM_ERR_ARG;
    }
    if (T && (ld < h->M || !x || !scores)) {
        h->setError(std::string(who) + ": invalid argument");
        return GMM_ERR_ARG;
    }
    return GMM_OK;
}

int gmm_score_host(gmm_handle* h, uint32_t T, const float* x, float* scores, uint32_t* best, size_t ld) {
    if (!h)
        return GMM_ERR_ARG;
    int rc = checkScoreHostArgs(h, T, x, scores, ld, "gmm_score_host");
    if (rc || T == 0)
        return rc;
    cudaSetDevice(h->device);
    cudaStream_t done = nullptr;
    if ((rc = scoreHostEnqueue(h, T, x, scores, best, ld, &done)))
        return rc;
    GMM_CUDA_TRY(h, cudaStreamSynchronize(done));
    return GMM_OK;
}

int gmm_score_async(gmm_handle* h, uint32_t T, const float* x, float* scores, uint32_t* best, size_t ld,
                    uint64_t* token) {
    if (!h || !token)
        return GMM_ERR_ARG;
    int rc = checkScoreHostArgs(h, T, x, scores, ld, "gmm_score_async");
    if (rc)
        return rc;
    cudaSetDevice(h->device);
    const int slot = (int)(h->asyncSeq % gmm_handle::kAsyncLanes);
    if (!h->asyncEvent[slot])
        GMM_CUDA_TRY(h, cudaEventCreateWithFlags(&h->asyncEvent[slot], cudaEventDisableTiming));
    if (h->asyncToken[slot])  // the lane's previous call still owns its scratch
        GMM_CUDA_TRY(h, cudaEventSynchronize(h->asyncEvent[slot]));
    const int back = h->lane;
    if ((rc = h->useLane(1 + slot)))
        return rc;
    cudaStream_t done = h->stream;
    if (T)
        rc = scoreHostEnqueue(h, T, x, scores, best, ld, &done);
    cudaError_t e = rc ? cudaSuccess : cudaEventRecord(h->asyncEvent[slot], done);
    h->useLane(back);
    if (rc)
        return rc;
    GMM_CUDA_TRY(h, e);
    h->asyncToken[slot] = ++h->asyncSeq;
    *token              = h->asyncToken[slot];
    return GMM_OK;
}

int gmm_wait(gmm_handle* h, uint64_t token) {
    if (!h)
        return GMM_ERR_ARG;
    if (token == 0 || token > h->asyncSeq) {
        h->setError("gmm_wait: unknown token");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    for (int i = 0; i < gmm_handle::kAsyncLanes; ++i)
        if (h->asyncToken[i] == token)
            GMM_CUDA_TRY(h, cudaEventSynchronize(h->asyncEvent[i]));
    return GMM_OK;  // an older token: its lane has been reused, which waited for it
}

int gmm_align_device(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                     const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                     const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                     const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState, float* pathScore) {
    if (!h)
        return GMM_ERR_ARG;
    if (nSeg == 0)
        return GMM_OK;
    if (!frameOffsets || !stateOffsets || !arcOffsets || !initialState || !finalWeight || !pathEmission || !pathScore ||
        (frameOffsets[nSeg] && !d_scores) || (arcOffsets[stateOffsets[nSeg]] && (!arcTarget || !arcEmission || !arcWeight))) {
        h->setError("gmm_align: NULL argument");
        return GMM_ERR_ARG;
    }
    if (int rc = gmm_gather_check(h, "gmm_align_device"))
        return rc;
    cudaSetDevice(h->device);
    return gmm_align_launch(h, nSeg, frameOffsets, d_scores, ld, stateOffsets, arcOffsets, arcTarget, arcEmission,
                            arcWeight, initialState, finalWeight, pathEmission, pathState, pathScore);
}

uint64_t gmm_align_ordered_count(const gmm_handle* h) {
    return h ? h->alignOrdered : 0;
}

int gmm_set_consumer_score_scale(gmm_handle* h, float scale) {
    if (!h)
        return GMM_ERR_ARG;
    if (!(scale > 0.f) || std::isinf(scale)) {
        h->setError("gmm_set_consumer_score_scale: the scale must be positive and finite");
        return GMM_ERR_ARG;
    }
    h->consumerScale = scale;
    return GMM_OK;
}

}  // extern "C"

// One group of whole segments: upload, score (compact block when that saves tensor work), align, accumulate --
// everything enqueued on the active lane, nothing waited for; `fin` completes it (gmm_align_finish).
// Baum-Welch groups (bw != null): posterior rows instead of a best path; rows go to bw->weights (host, may be NULL)
// and, with `accumulate`, into the estimators with the posterior as weight.
struct BwRequest {
    float*  weights;  // host [frames x ldW] or NULL
    size_t  ldW;
    double* total;    // host [nSeg] or NULL
};

// Callers whose arcs carry a LABEL that is not the emission index (RASR: allophone states, several of which may be
// tied to one mixture): arcEmission then holds segment-local label slots, and mixture[offsets[s] + slot] is the
// mixture that scores the label.  The label lists take the place of the emission lists: one column per label.
struct LabelMap {
    const uint32_t* offsets;  // [nSeg + 1]
    const uint32_t* mixture;
};

static int scoreAlignGroup(gmm_handle* h, bool accumulate, uint32_t s0, uint32_t s1, const uint32_t* frameOffsets,
                           const float* x, const uint32_t* stateOffsets, const uint32_t* arcOffsets,
                           const uint32_t* arcTarget, const uint32_t* arcEmission, const float* arcWeight,
                           const uint32_t* initialState, const float* finalWeight, uint32_t* pathEmission,
                           uint32_t* pathState, float* pathScore, AlignDeferred* fin, const BwRequest* bw = nullptr,
                           const LabelMap* labels = nullptr) {
    const size_t   devLd = h->M;
    const uint32_t f0 = frameOffsets[s0], T = frameOffsets[s1] - f0, n = s1 - s0;
    static const bool trace = getenv("GMM_TRACE") != nullptr;
    auto              clk   = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double            tr[6] = {clk(), 0, 0, 0, 0, 0};
    std::vector<uint32_t> fo(n + 1), so(n + 1);
    for (uint32_t i = 0; i <= n; ++i) {
        fo[i] = frameOffsets[s0 + i] - f0;
        so[i] = stateOffsets[s0 + i] - stateOffsets[s0];
    }
    std::vector<uint32_t> ao(so[n] + 1);
    const uint32_t        st0 = stateOffsets[s0], a0 = arcOffsets[st0];
    for (uint32_t s = 0; s <= so[n]; ++s)
        ao[s] = arcOffsets[st0 + s] - a0;
    void *dx = nullptr, *ds = nullptr;
    const uint32_t* dPath = nullptr;
    int   rc;
    const uint32_t inDim = h->lda.valid ? h->lda.baseDim : h->D;
    traceMark(h, (int)s0, "group start");
    if ((rc = uploadFrames(h, h->D, n, fo.data(), x + (size_t)f0 * inDim, &dx)))
        return rc;
    traceMark(h, (int)s0, "features on device");
    tr[1] = clk();
    // Lazy evaluation as the reference's cached context scorers do it (FeatureScorer.hh:170-230): only the
    // emissions a segment's automaton can ask for are scored, into a compact block whose column j is entry j of
    // the segment's emission list; the bits are those of the full pass.  Taken when it saves tensor work.
    AlignLists             lists;
    std::vector<uint32_t>& emOff = lists.emOff;
    std::vector<uint32_t>& emList = lists.emList;
    bool                  compact = false, haveLists = false;
    size_t                ldUse   = devLd;
    const char*           env     = getenv("GMM_ALIGN_SUBSET");
    std::vector<uint32_t> arcMixture;  // label callers: the mixture behind every arc's label
    if (labels) {
        const uint32_t l0 = labels->offsets[s0], nGroupArcs = ao[so[n]];
        if (nGroupArcs && !arcEmission) {
            h->setError("gmm_score_align_labels: arcLabel is NULL");
            return GMM_ERR_ARG;
        }
        emOff.resize(n + 1);
        for (uint32_t i = 0; i <= n; ++i) {
            if (labels->offsets[s0 + i] < (i ? labels->offsets[s0 + i - 1] : l0)) {
                h->setError("gmm_score_align_labels: label offsets not ascending");
                return GMM_ERR_ARG;
            }
            emOff[i] = labels->offsets[s0 + i] - l0;
        }
        emList.assign(labels->mixture + l0, labels->mixture + l0 + emOff[n]);
        for (uint32_t m : emList)
            if (m >= h->M) {
                h->setError("gmm_score_align_labels: label mapped to a mixture >= M");
                return GMM_ERR_ARG;
            }
        lists.arcSlot.assign(nGroupArcs ? arcEmission + a0 : nullptr, nGroupArcs ? arcEmission + a0 + nGroupArcs : nullptr);
        arcMixture.resize(nGroupArcs);
        for (uint32_t i = 0; i < n; ++i) {
            const uint32_t len = emOff[i + 1] - emOff[i];
            for (uint32_t a = ao[so[i]]; a < ao[so[i + 1]]; ++a) {
                if (lists.arcSlot[a] >= len) {
                    h->setError("gmm_score_align_labels: arc label outside its segment's label list");
                    return GMM_ERR_ARG;
                }
                arcMixture[a] = emList[emOff[i] + lists.arcSlot[a]];
            }
        }
        haveLists = true;
    }
    const uint32_t* arcEm = labels ? arcMixture.data() : arcEmission ? arcEmission + a0 : nullptr;
    if (T && gmm_tc_subset_capable(h) && !(env && env[0] == '0') && arcEmission &&
        (haveLists || gmm_align_emission_lists(h, n, so.data(), ao.data(), arcEm, h->M, lists))) {
        haveLists = true;
        const uint32_t mpt = h->tensor.mixPerTile;
        uint32_t       maxList = 0;
        double         subsetWork = 0, fullWork = (double)((T + 127) / 128) * ((h->M + mpt - 1) / mpt);
        for (uint32_t i = 0; i < n; ++i) {
            const uint32_t len = emOff[i + 1] - emOff[i];
            maxList            = std::max(maxList, len);
            subsetWork += (double)((fo[i + 1] - fo[i] + 127) / 128) * ((len + mpt - 1) / mpt);
        }
        compact = (env && env[0] == '1') || subsetWork < 0.6 * fullWork;
        if (compact)
            ldUse = gmm_tc_subset_ld(h, std::max(1u, maxList));
    }
    if ((rc = h->reserve(SCR_SCORE, (size_t)std::max(1u, T) * ldUse * 4, &ds)))
        return rc;
    tr[2] = clk();
    if (T) {
        if (compact)
            rc = gmm_tc_score_subset(h, n, fo.data(), emOff.data(), emList.data(), (const float*)dx, (float*)ds, ldUse);
        else
            rc = gmm_score_device(h, T, (const float*)dx, (float*)ds, nullptr, devLd);
        if (rc)
            return rc;
    }
    if (T && h->consumerScale != 1.f &&
        (rc = gmm_scale_block(h, (float*)ds, T, compact ? ldUse : (size_t)h->M, ldUse, h->consumerScale)))
        return rc;
    traceMark(h, (int)s0, "scored");
    tr[3] = clk();
    if (bw) {
        if (!haveLists && !gmm_align_emission_lists(h, n, so.data(), ao.data(), arcEm, h->M, lists)) {
            h->setError("gmm_score_baum_welch: arc offsets not ascending or emission index >= M");
            return GMM_ERR_ARG;
        }
        uint32_t maxList = 1;
        for (uint32_t i = 0; i < n; ++i)
            maxList = std::max(maxList, lists.emOff[i + 1] - lists.emOff[i]);
        if (bw->weights && bw->ldW < maxList) {
            h->setError("gmm_score_baum_welch: weights block narrower than a segment's emission list");
            return GMM_ERR_ARG;
        }
        const size_t ldP = (maxList + 31) & ~(size_t)31;
        void*        dp  = nullptr;
        BwDevice     dev;
        if ((rc = h->reserve(SCR_POST, (size_t)std::max(1u, T) * ldP * 4, &dp)))
            return rc;
        if ((rc = gmm_bw_launch(h, n, fo.data(), (const float*)ds, ldUse, so.data(), ao.data(),
                                arcTarget ? arcTarget + a0 : nullptr, arcEm,
                                arcWeight ? arcWeight + a0 : nullptr, initialState + s0, finalWeight + st0, (float*)dp, ldP,
                                nullptr, compact, h->M, &lists, false, &dev)))
            return rc;
        if (accumulate && T &&
            (rc = gmm_accum_rows_launch(h, n, T, (const float*)dx, dev.frameOff, dev.emOff, dev.emList, (const float*)dp, ldP)))
            return rc;
        fin->pending  = true;
        fin->nSeg     = n;
        fin->dWeights = (const float*)dp;
        fin->weights  = bw->weights ? bw->weights + (size_t)f0 * bw->ldW : nullptr;
        fin->ldWdev   = ldP;
        fin->ldWhost  = bw->ldW;
        fin->wRows    = T;
        fin->wCols    = bw->weights ? (uint32_t)std::min<size_t>(bw->ldW, ldP) : 0;
        fin->dTotal   = dev.total;
        fin->total    = bw->total ? bw->total + s0 : nullptr;
        return GMM_OK;
    }
    if ((rc = gmm_align_launch(h, n, fo.data(), (const float*)ds, ldUse, so.data(), ao.data(),
                               arcTarget ? arcTarget + a0 : nullptr, arcEm,
                               arcWeight ? arcWeight + a0 : nullptr, initialState + s0, finalWeight + st0,
                               pathEmission ? pathEmission + f0 : nullptr, pathState ? pathState + f0 : nullptr,
                               pathScore ? pathScore + s0 : nullptr, &dPath, compact, h->M, fin,
                               haveLists ? &lists : nullptr, labels != nullptr)))
        return rc;
    traceMark(h, (int)s0, "aligned");
    // the aligned frames go straight into the estimators: the path is the per-frame mixture (kNoArc: skip)
    tr[4] = clk();
    if (accumulate && T && (rc = gmm_accum_launch(h, T, (const float*)dx, dPath, nullptr, nullptr)))
        return rc;
    if (trace)
        fprintf(stderr, "[gmm] group of %u segments, host ms: offsets+upload %.3f, lists %.3f, score enqueue %.3f, align enqueue %.3f, accumulate %.3f\n",
                n, tr[1] - tr[0], tr[2] - tr[1], tr[3] - tr[2], tr[4] - tr[3], clk() - tr[4]);
    return GMM_OK;
}

static int scoreAlign(gmm_handle* h, bool accumulate, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                         const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                         const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                         const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState, float* pathScore,
                         const BwRequest* bw = nullptr, const LabelMap* labels = nullptr) {
    if (!h)
        return GMM_ERR_ARG;
    if (labels && (!labels->offsets || (nSeg && labels->offsets[nSeg] > labels->offsets[0] && !labels->mixture))) {
        h->setError("gmm_score_align_labels: NULL label list");
        return GMM_ERR_ARG;
    }
    if (!h->floatValid) {
        h->setError("gmm_score_align: no float model uploaded");
        return GMM_ERR_ARG;
    }
    if (nSeg == 0)
        return GMM_OK;
    if (accumulate && !h->accum.valid) {
        h->setError("gmm_score_align_accumulate: no accumulators (gmm_accumulator_create)");
        return GMM_ERR_ARG;
    }
    if (!frameOffsets || !stateOffsets || !arcOffsets || !initialState || !finalWeight ||
        (!accumulate && !bw && (!pathEmission || !pathScore)) || (frameOffsets[nSeg] && !x)) {
        h->setError("gmm_score_align: NULL argument");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    // Groups of whole segments, alternating between the two lanes of the handle: while group g is scored and aligned
    // on one stream, group g+1's tables are built on the host and its features cross PCIe on the other.  A group is
    // large enough to fill the alignment kernel (one CTA per segment, two per SM), and bounded so that its score
    // block (which never leaves the device) stays within 10 GB.
    const size_t   maxFrames = std::max<size_t>(1, ((size_t)10 << 30) / ((size_t)h->M * 4));
    const char*    envG      = getenv("GMM_ALIGN_GROUP");
    const uint32_t groupSegs = envG ? (uint32_t)std::max(1, atoi(envG)) : (uint32_t)h->smCount * 2;
    // three lanes by default: with two, a lane's next upload cannot be enqueued before its previous group is complete
    // (the group's frames are read until its last kernel), and PCIe idles for that long
    const char*    envL      = getenv("GMM_LANES");
    const int      nLanes    = nSeg > groupSegs ? std::min(gmm_handle::kMaxLanes, std::max(2, envL ? atoi(envL) : 3)) : 1;
    const bool     twoLanes  = nLanes > 1;
    AlignDeferred  fin[gmm_handle::kMaxLanes];
    int            rc = GMM_OK, lane = 0;
    if (getenv("GMM_TRACE")) {
        cudaEventCreate(&g_traceStart);
        cudaEventRecord(g_traceStart, h->stream);
    }
    if (twoLanes) {  // the other lanes start after whatever the handle's stream holds so far (model upload, ...)
        for (int l = nLanes - 1; l >= 0; --l)
            if ((rc = h->useLane(l)))
                return rc;
        GMM_CUDA_TRY(h, cudaEventRecord(h->laneEvent, h->stream));
        for (int l = 1; l < nLanes; ++l)
            GMM_CUDA_TRY(h, cudaStreamWaitEvent(h->parked[l].stream, h->laneEvent, 0));
    }
    uint32_t s0 = 0;
    while (s0 < nSeg && rc == GMM_OK) {
        uint32_t s1 = s0 + 1;
        while (s1 < nSeg && s1 - s0 < groupSegs && (size_t)(frameOffsets[s1 + 1] - frameOffsets[s0]) <= maxFrames)
            ++s1;
        if (twoLanes)
            rc = h->useLane(lane);
        const auto tw0 = std::chrono::steady_clock::now();
        if (rc == GMM_OK)
            rc = gmm_align_finish(h, fin[lane]);  // the lane's previous group: its scratch is about to be reused
        if (getenv("GMM_TRACE"))
            fprintf(stderr, "[gmm] lane %d: waited %.3f ms for its previous group\n", lane,
                    std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tw0).count());
        if (rc == GMM_OK)
            rc = scoreAlignGroup(h, accumulate, s0, s1, frameOffsets, x, stateOffsets, arcOffsets, arcTarget, arcEmission,
                                 arcWeight, initialState, finalWeight, pathEmission, pathState, pathScore, &fin[lane], bw,
                                 labels);
        if (twoLanes)
            lane = (lane + 1) % nLanes;
        s0 = s1;
    }
    for (int k = 0; k < nLanes; ++k) {  // oldest group first; on an error still drain every lane
        const int l = (lane + k) % nLanes;
        if (twoLanes)
            h->useLane(l);
        const int r = gmm_align_finish(h, fin[l]);
        if (twoLanes)
            cudaStreamSynchronize(h->stream);
        if (rc == GMM_OK)
            rc = r;
    }
    h->useLane(0);
    if (g_traceStart) {
        for (const TraceMark& m : g_trace) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, g_traceStart, m.ev);
            fprintf(stderr, "[gmm-timeline] %8.3f ms  lane %d  segment %5d  %s\n", ms, m.lane, m.group, m.what);
            cudaEventDestroy(m.ev);
        }
        g_trace.clear();
        cudaEventDestroy(g_traceStart);
        g_traceStart = nullptr;
    }
    return rc;
}

extern "C" {

int gmm_score_align_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                         const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                         const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                         const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState, float* pathScore) {
    return scoreAlign(h, false, nSeg, frameOffsets, x, stateOffsets, arcOffsets, arcTarget, arcEmission, arcWeight,
                      initialState, finalWeight, pathEmission, pathState, pathScore);
}

int gmm_score_align_accumulate_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                                    const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                                    const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                                    const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState,
                                    float* pathScore) {
    return scoreAlign(h, true, nSeg, frameOffsets, x, stateOffsets, arcOffsets, arcTarget, arcEmission, arcWeight,
                      initialState, finalWeight, pathEmission, pathState, pathScore);
}

int gmm_score_align_labels_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                                const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                                const uint32_t* arcLabel, const float* arcWeight, const uint32_t* initialState,
                                const float* finalWeight, const uint32_t* labelOffsets, const uint32_t* labelMixture,
                                uint32_t* pathLabel, uint32_t* pathState, float* pathScore) {
    const LabelMap labels = {labelOffsets, labelMixture};
    return scoreAlign(h, false, nSeg, frameOffsets, x, stateOffsets, arcOffsets, arcTarget, arcLabel, arcWeight,
                      initialState, finalWeight, pathLabel, pathState, pathScore, nullptr, &labels);
}

// ---- Baum-Welch alignment ---------------------------------------------------------------------------------------
int gmm_emission_lists(gmm_handle* h, uint32_t nSeg, const uint32_t* stateOffsets, const uint32_t* arcOffsets,
                       const uint32_t* arcEmission, uint32_t nEmissions, uint32_t* listOffsets, uint32_t* list) {
    if (!h)
        return GMM_ERR_ARG;
    if (!stateOffsets || !arcOffsets || !listOffsets || (nSeg && arcOffsets[stateOffsets[nSeg]] && (!arcEmission || !list))) {
        h->setError("gmm_emission_lists: NULL argument");
        return GMM_ERR_ARG;
    }
    AlignLists L;
    if (!gmm_align_emission_lists(h, nSeg, stateOffsets, arcOffsets, arcEmission, nEmissions, L)) {
        h->setError("gmm_emission_lists: arc offsets not ascending or emission index >= nEmissions");
        return GMM_ERR_ARG;
    }
    memcpy(listOffsets, L.emOff.data(), (size_t)(nSeg + 1) * 4);
    if (!L.emList.empty())
        memcpy(list, L.emList.data(), L.emList.size() * 4);
    return GMM_OK;
}

int gmm_baum_welch_device(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                          const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                          const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                          const float* finalWeight, float* d_weights, size_t ldW, double* totalScore) {
    if (!h)
        return GMM_ERR_ARG;
    if (nSeg == 0)
        return GMM_OK;
    if (!frameOffsets || !stateOffsets || !arcOffsets || !initialState || !finalWeight ||
        (frameOffsets[nSeg] && (!d_scores || !d_weights)) ||
        (arcOffsets[stateOffsets[nSeg]] && (!arcTarget || !arcEmission || !arcWeight))) {
        h->setError("gmm_baum_welch: NULL argument");
        return GMM_ERR_ARG;
    }
    if (int rc = gmm_gather_check(h, "gmm_baum_welch_device"))
        return rc;
    cudaSetDevice(h->device);
    return gmm_bw_launch(h, nSeg, frameOffsets, d_scores, ld, stateOffsets, arcOffsets, arcTarget, arcEmission, arcWeight,
                         initialState, finalWeight, d_weights, ldW, totalScore, false, 0, nullptr, true);
}

int gmm_score_baum_welch_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                              const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                              const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                              const float* finalWeight, float* weights, size_t ldW, double* totalScore) {
    if (h && nSeg && frameOffsets && frameOffsets[nSeg] && !weights) {
        h->setError("gmm_score_baum_welch: weights is NULL");
        return GMM_ERR_ARG;
    }
    const BwRequest bw = {weights, ldW, totalScore};
    return scoreAlign(h, false, nSeg, frameOffsets, x, stateOffsets, arcOffsets, arcTarget, arcEmission, arcWeight,
                      initialState, finalWeight, nullptr, nullptr, nullptr, &bw);
}

int gmm_score_baum_welch_labels_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                                     const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                                     const uint32_t* arcLabel, const float* arcWeight, const uint32_t* initialState,
                                     const float* finalWeight, const uint32_t* labelOffsets,
                                     const uint32_t* labelMixture, float* weights, size_t ldW, double* totalScore) {
    if (h && nSeg && frameOffsets && frameOffsets[nSeg] && !weights) {
        h->setError("gmm_score_baum_welch_labels: weights is NULL");
        return GMM_ERR_ARG;
    }
    const BwRequest bw     = {weights, ldW, totalScore};
    const LabelMap  labels = {labelOffsets, labelMixture};
    return scoreAlign(h, false, nSeg, frameOffsets, x, stateOffsets, arcOffsets, arcTarget, arcLabel, arcWeight,
                      initialState, finalWeight, nullptr, nullptr, nullptr, &bw, &labels);
}

int gmm_score_baum_welch_accumulate_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                                         const uint32_t* stateOffsets, const uint32_t* arcOffsets,
                                         const uint32_t* arcTarget, const uint32_t* arcEmission, const float* arcWeight,
                                         const uint32_t* initialState, const float* finalWeight, double* totalScore) {
    const BwRequest bw = {nullptr, 0, totalScore};
    return scoreAlign(h, true, nSeg, frameOffsets, x, stateOffsets, arcOffsets, arcTarget, arcEmission, arcWeight,
                      initialState, finalWeight, nullptr, nullptr, nullptr, &bw);
}

// ---- front end on the device -------------------------------------------------------------------------------------
int gmm_set_lda(gmm_handle* h, uint32_t baseDim, uint32_t length, uint32_t right, uint32_t D, const float* matrix) {
    if (!h)
        return GMM_ERR_ARG;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    if (h->lda.matrix)
        cudaFree(h->lda.matrix);
    h->lda = LdaModel();
    if (!matrix || baseDim == 0)
        return GMM_OK;  // cleared: host entry points take model-dimension features again
    if (length == 0 || right >= length || D == 0) {
        h->setError("gmm_set_lda: need length >= 1, right < length, D >= 1");
        return GMM_ERR_ARG;
    }
    const size_t bytes = (size_t)D * baseDim * length * 4;
    GMM_CUDA_TRY(h, cudaMalloc((void**)&h->lda.matrix, bytes));
    GMM_CUDA_TRY(h, cudaMemcpyAsync(h->lda.matrix, matrix, bytes, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    h->lda.baseDim = baseDim;
    h->lda.length  = length;
    h->lda.right   = right;
    h->lda.D       = D;
    h->lda.valid   = true;
    return GMM_OK;
}

int gmm_lda_device(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_base, float* d_x) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->lda.valid) {
        h->setError("gmm_lda_device: no front end set (gmm_set_lda)");
        return GMM_ERR_ARG;
    }
    if (nSeg == 0)
        return GMM_OK;
    if (!frameOffsets || (frameOffsets[nSeg] > frameOffsets[0] && (!d_base || !d_x))) {
        h->setError("gmm_lda_device: NULL argument");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    return gmm_lda_launch(h, nSeg, frameOffsets, d_base, d_x);
}

// ---- training E-step -------------------------------------------------------------------------------------------
int gmm_accumulator_create(gmm_handle* h, uint32_t nMeans, const uint32_t* meanIdx, int viterbi, double weightThreshold) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->floatValid) {
        h->setError("gmm_accumulator_create: no float model uploaded");
        return GMM_ERR_ARG;
    }
    if (!meanIdx || nMeans == 0) {
        h->setError("gmm_accumulator_create: NULL mean index table");
        return GMM_ERR_ARG;
    }
    if (!viterbi && h->mode != GMM_MODE_SUM) {
        // CachedAssigningFeatureScorer::calculateDensityPosteriorProbabilities is only implemented by the sum scorer
        h->setError("gmm_accumulator_create: posterior-weighted accumulation needs the sum scorer (GMM_MODE_SUM)");
        return GMM_ERR_UNSUPPORTED;
    }
    for (uint32_t g = 0; g < h->G; ++g)
        if (meanIdx[g] >= nMeans) {
            h->setError("gmm_accumulator_create: mean index out of range");
            return GMM_ERR_ARG;
        }
    cudaSetDevice(h->device);
    AccumModel& am = h->accum;
    if (am.valid) {
        h->setError("gmm_accumulator_create: this float model already has accumulators (upload the model again to replace)");
        return GMM_ERR_ARG;
    }
    int rc;
    if ((rc = h->upload((void**)&am.meanIdx, meanIdx, (size_t)h->G * 4)))
        return rc;
    const size_t D = h->D, n = (size_t)nMeans * D + nMeans + (size_t)h->C * D + h->C + h->G;
    if ((rc = h->devAlloc((void**)&am.block, n * sizeof(double))))
        return rc;
    am.blockBytes = n * sizeof(double);
    am.meanSum    = am.block;
    am.meanW      = am.meanSum + (size_t)nMeans * D;
    am.covSum     = am.meanW + nMeans;
    am.covW       = am.covSum + (size_t)h->C * D;
    am.mixW       = am.covW + h->C;
    am.nMeans     = nMeans;
    am.viterbi    = viterbi ? 1 : 0;
    am.weightThreshold = weightThreshold;
    GMM_CUDA_TRY(h, cudaMemsetAsync(am.block, 0, am.blockBytes, h->stream));
    am.valid = true;
    return GMM_OK;
}

int gmm_accumulator_reset(gmm_handle* h) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->accum.valid) {
        h->setError("gmm_accumulator_reset: no accumulators");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    GMM_CUDA_TRY(h, cudaMemsetAsync(h->accum.block, 0, h->accum.blockBytes, h->stream));
    return GMM_OK;
}

int gmm_accumulator_read(gmm_handle* h, double* meanSum, double* meanWeight, double* covSum, double* covWeight,
                         double* mixWeight) {
    if (!h)
        return GMM_ERR_ARG;
    const AccumModel& am = h->accum;
    if (!am.valid) {
        h->setError("gmm_accumulator_read: no accumulators");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    const size_t D = h->D;
    auto down = [&](double* dst, const double* src, size_t n) {
        return dst && n ? cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream) : cudaSuccess;
    };
    GMM_CUDA_TRY(h, down(meanSum, am.meanSum, (size_t)am.nMeans * D));
    GMM_CUDA_TRY(h, down(meanWeight, am.meanW, am.nMeans));
    GMM_CUDA_TRY(h, down(covSum, am.covSum, (size_t)h->C * D));
    GMM_CUDA_TRY(h, down(covWeight, am.covW, h->C));
    GMM_CUDA_TRY(h, down(mixWeight, am.mixW, h->G));
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GMM_OK;
}

size_t gmm_accumulator_size(const gmm_handle* h) {
    return h && h->accum.valid ? h->accum.blockBytes / sizeof(double) : 0;
}

int gmm_accumulator_export_device(gmm_handle* h, double* d_dst) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->accum.valid || !d_dst) {
        h->setError("gmm_accumulator_export_device: no accumulators or NULL destination");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    GMM_CUDA_TRY(h, cudaMemcpyAsync(d_dst, h->accum.block, h->accum.blockBytes, cudaMemcpyDeviceToDevice, h->stream));
    return GMM_OK;
}

int gmm_accumulator_import_device(gmm_handle* h, const double* d_src) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->accum.valid || !d_src) {
        h->setError("gmm_accumulator_import_device: no accumulators or NULL source");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    GMM_CUDA_TRY(h, cudaMemcpyAsync(h->accum.block, d_src, h->accum.blockBytes, cudaMemcpyDeviceToDevice, h->stream));
    return GMM_OK;
}

int gmm_accumulate_device(gmm_handle* h, uint32_t T, const float* d_x, const uint32_t* d_mixture, const double* d_weight,
                          uint32_t* d_best) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->floatValid || !h->accum.valid) {
        h->setError("gmm_accumulate: no float model or no accumulators (gmm_accumulator_create)");
        return GMM_ERR_ARG;
    }
    if (T == 0)
        return GMM_OK;
    if (!d_x || !d_mixture) {
        h->setError("gmm_accumulate: NULL argument");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    return gmm_accum_launch(h, T, d_x, d_mixture, d_weight, d_best);
}

int gmm_accumulate_host(gmm_handle* h, uint32_t T, const float* x, const uint32_t* mixture, const double* weight,
                        uint32_t* best) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->floatValid || !h->accum.valid) {
        h->setError("gmm_accumulate: no float model or no accumulators (gmm_accumulator_create)");
        return GMM_ERR_ARG;
    }
    if (T == 0)
        return GMM_OK;
    if (!x || !mixture) {
        h->setError("gmm_accumulate: NULL argument");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    for (uint32_t t = 0; t < T; ++t)
        if (mixture[t] != 0xFFFFFFFFu && mixture[t] >= h->M) {
            h->setError("gmm_accumulate: mixture index out of range");
            return GMM_ERR_ARG;
        }
    void *dx = nullptr, *dm = nullptr;
    int   rc;
    const uint32_t one[2] = {0, T};
    if ((rc = uploadFrames(h, h->D, 1, one, x, &dx)))
        return rc;
    // mixture indices, weights and chosen densities share one scratch reservation
    const size_t oW = ((size_t)T * 4 + 255) & ~(size_t)255, oB = oW + (((size_t)T * 8 + 255) & ~(size_t)255);
    if ((rc = h->reserve(SCR_BEST, oB + (size_t)T * 4, &dm)))
        return rc;
    char* base = (char*)dm;
    GMM_CUDA_TRY(h, cudaMemcpyAsync(base, mixture, (size_t)T * 4, cudaMemcpyHostToDevice, h->stream));
    if (weight)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(base + oW, weight, (size_t)T * 8, cudaMemcpyHostToDevice, h->stream));
    if ((rc = gmm_accum_launch(h, T, (const float*)dx, (const uint32_t*)base, weight ? (const double*)(base + oW) : nullptr,
                               best ? (uint32_t*)(base + oB) : nullptr)))
        return rc;
    if (best)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(best, base + oB, (size_t)T * 4, cudaMemcpyDeviceToHost, h->stream));
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GMM_OK;
}

int gmm_posteriors_device(gmm_handle* h, uint32_t T, const float* d_x, float* d_post) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->floatValid || (T && (!d_x || !d_post))) {
        h->setError("gmm_posteriors: no float model uploaded or NULL buffer");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    return gmm_direct_posteriors(h, T, d_x, d_post);
}

int gmm_mixture_posteriors_host(gmm_handle* h, uint32_t T, const float* x, const uint32_t* mixture, float* post,
                                size_t ldP, float* score) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->floatValid || h->mode != GMM_MODE_SUM) {
        h->setError("gmm_mixture_posteriors_host: needs a float model uploaded with GMM_MODE_SUM (only the sum scorer "
                    "has density posteriors, GaussDiagonalMaximumFeatureScorer.cc:291-298)");
        return GMM_ERR_ARG;
    }
    if (T == 0)
        return GMM_OK;
    if (!x || !mixture || !post || ldP < h->maxK || (size_t)h->maxK * 4 > 48 * 1024) {
        h->setError("gmm_mixture_posteriors_host: invalid argument (ldP below the largest mixture?)");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    void *dx = nullptr, *dm = nullptr, *dp = nullptr, *ds = nullptr;
    int   rc;
    const uint32_t one[2] = {0, T};
    if ((rc = uploadFrames(h, h->D, 1, one, x, &dx)))
        return rc;
    if ((rc = h->reserve(SCR_MISC, (size_t)T * 4, &dm)))
        return rc;
    if ((rc = h->reserve(SCR_POST, (size_t)T * h->maxK * 4, &dp)))
        return rc;
    if (score && (rc = h->reserve(SCR_ROW, (size_t)T * 4, &ds)))
        return rc;
    GMM_CUDA_TRY(h, cudaMemcpyAsync(dm, mixture, (size_t)T * 4, cudaMemcpyHostToDevice, h->stream));
    GMM_CUDA_TRY(h, cudaMemsetAsync(dp, 0, (size_t)T * h->maxK * 4, h->stream));
    if ((rc = gmm_direct_mixture_posteriors(h, T, (const float*)dx, (const uint32_t*)dm, (float*)dp, h->maxK, (float*)ds,
                                            h->maxK)))
        return rc;
    GMM_CUDA_TRY(h, cudaMemcpy2DAsync(post, ldP * 4, dp, (size_t)h->maxK * 4, (size_t)h->maxK * 4, T,
                                      cudaMemcpyDeviceToHost, h->stream));
    if (score)
        GMM_CUDA_TRY(h, cudaMemcpyAsync(score, ds, (size_t)T * 4, cudaMemcpyDeviceToHost, h->stream));
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return GMM_OK;
}

int gmm_upload_quant(gmm_handle* h, uint32_t M, uint32_t G, uint32_t D, uint32_t Dpad, uint32_t C,
                     const uint32_t* offsets, const uint8_t* qmean, const uint32_t* cov_idx, const int32_t* qconst,
                     const float* isr_scaled, float scaling_squared) {
    if (!h)
        return GMM_ERR_ARG;
    cudaSetDevice(h->device);
    if (D == 0 || C == 0 || Dpad < D || Dpad % 16 != 0 || Dpad > 128 || (G && (!qmean || !cov_idx || !qconst)) ||
        !isr_scaled) {
        h->setError("gmm_upload_quant: invalid argument (Dpad must be a multiple of 16, at most 128)");
        return GMM_ERR_ARG;
    }
    int rc = checkOffsets(h, M, G, offsets);
    if (rc)
        return rc;
    for (uint32_t g = 0; g < G; ++g)
        if (cov_idx[g] >= C) {
            h->setError("gmm_upload_quant: covariance index out of range");
            return GMM_ERR_ARG;
        }
    releaseQuantModel(h);
    QuantModel& qm = h->quant;
    qm.M = M, qm.G = G, qm.D = D, qm.Dpad = Dpad, qm.C = C;
    qm.scalingSquared = scaling_squared;
    qm.scoreRule      = GMM_QUANT_RULE_SIMD;
    // qconst2 = qconst + |qmean|^2 with wrapping adds (the kernel adds |qx|^2 - 2 a.b)
    std::vector<int32_t> c2(G);
    for (size_t g = 0; g < G; ++g) {
        uint32_t s = 0;
        for (uint32_t d = 0; d < Dpad; ++d) {
            uint32_t v = qmean[g * Dpad + d];
            s += v * v;
        }
        c2[g] = (int32_t)((uint32_t)qconst[g] + s);
    }
    if ((rc = h->upload((void**)&qm.qmean, qmean, (size_t)G * Dpad, true)))
        return rc;
    if ((rc = h->upload((void**)&qm.covIdx, cov_idx, (size_t)G * 4, true)))
        return rc;
    if ((rc = h->upload((void**)&qm.qconst2, c2.data(), (size_t)G * 4, true)))
        return rc;
    if ((rc = h->upload((void**)&qm.isrS, isr_scaled, (size_t)C * D * 4, true)))
        return rc;
    if ((rc = h->upload((void**)&qm.offsets, offsets, (size_t)(M + 1) * 4, true)))
        return rc;
    GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    qm.valid = true;
    // integer tensor path where the model allows it (pooled covariance, no wrap-around); else dp4a
    return gmm_qtc_pack(h, offsets, qmean, c2.data());
}

int gmm_quant_score_device(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best,
                           float* d_score, uint8_t* d_qx, size_t ld) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->quant.valid) {
        h->setError("gmm_quant_score: no quantized model uploaded");
        return GMM_ERR_ARG;
    }
    if (ld < h->quant.M || (T && !d_x)) {
        h->setError("gmm_quant_score: invalid argument (ld < M or NULL features)");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    return gmm_quant_launch(h, T, d_x, d_q, d_best, d_score, d_qx, ld);
}

int gmm_quant_score_host(gmm_handle* h, uint32_t T, const float* x, int32_t* q, uint32_t* best, float* score,
                         uint8_t* qx, size_t ld) {
    if (!h)
        return GMM_ERR_ARG;
    if (!h->quant.valid) {
        h->setError("gmm_quant_score: no quantized model uploaded");
        return GMM_ERR_ARG;
    }
    if (T == 0)
        return GMM_OK;
    const QuantModel& qm = h->quant;
    if (ld < qm.M || !x) {
        h->setError("gmm_quant_score_host: invalid argument");
        return GMM_ERR_ARG;
    }
    cudaSetDevice(h->device);
    // with a front end the whole block of base features is transformed first (windows span chunk boundaries)
    void* dxAll = nullptr;
    if (h->lda.valid) {
        const uint32_t one[2] = {0, T};
        int            rc     = uploadFrames(h, qm.D, 1, one, x, &dxAll);
        if (rc)
            return rc;
    }
    // bound device scratch: frames in chunks
    const size_t perFrame = (size_t)qm.M * 12 + (qx ? (size_t)qm.C * qm.Dpad : 0) + (size_t)qm.D * 4;
    uint32_t     chunk    = (uint32_t)std::max<size_t>(1, std::min<size_t>(T, ((size_t)1 << 30) / perFrame));
    for (uint32_t t0 = 0; t0 < T; t0 += chunk) {
        const uint32_t n = std::min(chunk, T - t0);
        void *dx = nullptr, *dq = nullptr, *db = nullptr, *ds = nullptr, *dqx = nullptr;
        int   rc;
        if (dxAll)
            dx = (float*)dxAll + (size_t)t0 * qm.D;
        else if ((rc = h->reserve(SCR_X, (size_t)n * qm.D * 4, &dx)))
            return rc;
        if (q && (rc = h->reserve(SCR_Q, (size_t)n * qm.M * 4, &dq)))
            return rc;
        if (best && (rc = h->reserve(SCR_BEST, (size_t)n * qm.M * 4, &db)))
            return rc;
        if (score && (rc = h->reserve(SCR_SCORE, (size_t)n * qm.M * 4, &ds)))
            return rc;
        if (qx && (rc = h->reserve(SCR_MISC, (size_t)n * qm.C * qm.Dpad, &dqx)))
            return rc;
        if (!dxAll)
            GMM_CUDA_TRY(h, cudaMemcpyAsync(dx, x + (size_t)t0 * qm.D, (size_t)n * qm.D * 4, cudaMemcpyHostToDevice,
                                            h->stream));
        if ((rc = gmm_quant_launch(h, n, (const float*)dx, (int32_t*)dq, (uint32_t*)db, (float*)ds, (uint8_t*)dqx,
                                   qm.M)))
            return rc;
        if (q && (rc = copyOut(h, q + (size_t)t0 * ld, dq, n, qm.M, ld, qm.M)))
            return rc;
        if (best && (rc = copyOut(h, best + (size_t)t0 * ld, db, n, qm.M, ld, qm.M)))
            return rc;
        if (score && (rc = copyOut(h, score + (size_t)t0 * ld, ds, n, qm.M, ld, qm.M)))
            return rc;
        if (qx)
            GMM_CUDA_TRY(h, cudaMemcpyAsync(qx + (size_t)t0 * qm.C * qm.Dpad, dqx, (size_t)n * qm.C * qm.Dpad,
                                            cudaMemcpyDeviceToHost, h->stream));
        GMM_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    }
    return GMM_OK;
}

}  // extern "C"
