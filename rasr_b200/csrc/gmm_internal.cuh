// Internal declarations shared by the translation units of libgmm_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <utility>
#include <vector>

#include "../../include/gmm_b200.h"
#include "gmm_host_pool.h"

#define GMM_CUDA_TRY(h, expr)                                                                          \
    do {                                                                                               \
        cudaError_t e__ = (expr);                                                                      \
        if (e__ != cudaSuccess) {                                                                      \
            (h)->setError(std::string(#expr) + ": " + cudaGetErrorString(e__));                        \
            return GMM_ERR_CUDA;                                                                       \
        }                                                                                              \
    } while (0)

// ---- float model as the direct kernel reads it (slot-major, padded to DP floats per row) ---------
struct DirectModel {
    uint32_t  DP      = 0;        // D rounded up to a multiple of 4
    float*    mean    = nullptr;  // [G x DP], zero padded
    float*    isr     = nullptr;  // [C x DP], zero padded
    uint32_t* covIdx  = nullptr;  // [G]
    float*    w2      = nullptr;  // [G] minus2LogWeight
    float*    logNorm = nullptr;  // [C] logNormalizationFactor
    uint32_t* offsets = nullptr;  // [M+1]
};

// ---- float model as the tensor kernel reads it (see gmm_tc.cu for the tile image) -----------------
struct TensorModel {
    bool      valid     = false;
    bool      pooled    = false;   // C == 1: whitened form, K = D; else K = 2D ([x^2, x] . [iv, -2 mu iv])
    uint32_t  kSteps    = 0;       // 16-wide k-steps per precision part
    uint32_t  nAtoms    = 0;       // swizzle atoms per operand row (both parts)
    uint32_t  nTiles    = 0;       // 256-column tiles of the packed B operand
    uint32_t  tileBytes = 0;       // bytes per B tile image
    uint8_t*  bTiles    = nullptr; // [nTiles x tileBytes]
    float*    center    = nullptr; // [D] translation applied to features and means (pooled form)
    float*    scale     = nullptr; // [D] per-dimension factor of the frame preparation (pooled: isr)
    uint32_t  partW     = 128;     // packed columns reduced by one epilogue warp (128: 8 warps, 64: 16 warps)
    uint32_t  maxMixPerPart = 0;
    uint32_t  swz       = 128;     // swizzle atom width in bytes (128 or 64)
    uint32_t  aTileBytes = 0;
    uint32_t  buffersB  = 0;       // B tile buffers that fit in shared memory (1 or 2)
    uint32_t  stagesA   = 0;       // A tile ring depth with the arg-min staging enabled
    // regular model: every mixture takes the same number of packed columns (mixCols, a divisor of partW) and none
    // is empty, so mixture m sits at rows (m % mixPerTile) * mixCols of tile m / mixPerTile and a tile over ANY
    // set of mixtures can be assembled from whole swizzle atoms of the canonical image (subset scoring)
    // GMM_TC_SPLIT=fp8 (experimental, pooled models): part 2 of the operand rows holds e4m3 bytes (lo parts times 2^8) -- per 32-byte
    // k-step 16 x lo and 16 x hi of the same dimensions (A: lo | hi, B: hi | lo) -- and ONE kind::f8f6f4 instruction
    // per k-step adds both correction products: 6 MMA instructions per tile instead of 9
    bool      fp8       = false;
    bool      accuracyOk = true;   // centred, whitened means within 8 sigma rms: the expanded form keeps 1e-4 relative
    bool      regular   = false;
    uint32_t  mixCols   = 0, mixPerTile = 0;
};

struct QuantModel {
    bool      valid   = false;
    uint32_t  M = 0, G = 0, D = 0, Dpad = 0, C = 0;
    float     scalingSquared = 0.f;
    int       scoreRule      = 0;    // GMM_QUANT_RULE_*: how q becomes the final f32 score
    // what the kernels take: s^2, negated when the final score is the batch-int scorer's f32 division q / (2 s^2)
    double finalScoreArg() const { return scoreRule ? -(double)scalingSquared : (double)scalingSquared; }
    uint8_t*  qmean   = nullptr;  // [G x Dpad]
    uint32_t* covIdx  = nullptr;  // [G]
    int32_t*  qconst2 = nullptr;  // [G] qconst + sum_d qmean^2 (wrapping)
    float*    isrS    = nullptr;  // [C x D]
    uint32_t* offsets = nullptr;  // [M+1]
};

// estimators of the training E-step (gmm_accum.cu): f64 sums per mean / covariance / (mixture, density)
struct AccumModel {
    bool      valid   = false;
    int       viterbi = 1;
    double    weightThreshold = 0.0;
    uint32_t  nMeans  = 0;
    uint32_t* meanIdx = nullptr;  // [G] mean of every (mixture, density) slot
    double*   block   = nullptr;  // one allocation: the five arrays below
    size_t    blockBytes = 0;
    double *  meanSum = nullptr, *meanW = nullptr, *covSum = nullptr, *covW = nullptr, *mixW = nullptr;
};

// front end applied on the device when frames arrive from the host (gmm_lda.cu)
struct LdaModel {
    bool     valid   = false;
    uint32_t baseDim = 0, length = 0, right = 0, D = 0;
    float*   matrix  = nullptr;  // [D x baseDim * length], row major
};

// quantized scorer on the integer tensor path (gmm_qtc.cu); valid only for models that meet its requirements
struct QTensorModel {
    bool      valid      = false;
    uint32_t  K          = 0;       // operand row bytes: Dpad + 16 or 32 spare bytes that carry the constants
    uint32_t  kSteps     = 0;       // 32-byte k-steps of the contraction (K / 32)
    uint32_t  nAtoms     = 0;       // 64-byte swizzle atoms per A operand row
    uint32_t  nAtomsB    = 0;       // ... per B operand row (one more k-step: the variant of the last)
    uint32_t  nTiles     = 0;       // 256-column tiles of the packed B operand
    uint32_t  tileBytes  = 0;       // bytes per B tile image (operand rows, Kc table, TileMeta)
    uint32_t  aTileBytes = 0;
    uint8_t*  bTiles     = nullptr;
    uint32_t  partW      = 128;     // packed columns reduced by one epilogue warp (64: 16 warps)
    uint32_t  maxMixPerPart = 0;
    uint32_t  buffersB   = 0, stagesA = 0;
    uint32_t  qBase      = 0;       // max c2 (wrapping)
    uint32_t  qAbsMax    = 0;       // bound on |q| over all frames (f32 path of the final score needs < 2^24)
};

struct gmm_handle {
    int          device     = 0;
    int          smCount    = 0;
    cudaStream_t stream     = nullptr;
    bool         ownStream  = false;
    uint64_t     launches   = 0;
    uint64_t     alignOrdered = 0;  // segments aligned by the ordered (tie-resolving) pass so far
    float        consumerScale = 1.f;  // the *_host alignment consumers read fl(scale * score) (gmm_set_consumer_score_scale)
    bool         timing     = false;                            // gmm_kernel_timing
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> timed;     // brackets of scoring-kernel launches
    void timeBegin() {
        if (!timing)
            return;
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        cudaEventRecord(a, stream);
        timed.push_back(std::make_pair(a, b));
    }
    void timeEnd() {
        if (timing && !timed.empty())
            cudaEventRecord(timed.back().second, stream);
    }
    std::string  error;

    // float engine
    bool        floatValid = false;
    uint32_t    M = 0, G = 0, D = 0, C = 0;
    uint32_t    maxK = 0;  // most densities in one mixture
    int         mode = GMM_MODE_MAX, kernel = GMM_KERNEL_DIRECT;
    DirectModel direct;
    TensorModel tensor;
    QuantModel  quant;
    QTensorModel qtensor;
    AccumModel   accum;
    LdaModel     lda;

    // scratch (grown on demand)
    void*  scratch[16]      = {};
    size_t scratchBytes[16] = {};
    // Lanes of the pipelined host entry points (gmm_score_align_*_host): each its own stream, scratch and staging, so
    // that one group of segments is uploaded and prepared while the previous ones are scored and aligned.  useLane()
    // makes a lane the one `stream` / `scratch` refer to; every kernel launcher keeps using those two names.
    struct LaneState {  // a parked lane: its stream, scratch set and pinned staging blocks
        cudaStream_t stream           = nullptr;
        void*        scratch[16]      = {};
        size_t       scratchBytes[16] = {};
        void*        pinned[2]        = {};
        size_t       pinnedBytes[2]   = {};
        cudaEvent_t  pinEvent[2]      = {};
    };
    static constexpr int kMaxLanes = 4;
    LaneState    parked[kMaxLanes];          // parked[i] holds lane i's state while another lane is active
    cudaEvent_t  laneEvent           = nullptr;
    int          lane                = 0;    // the active lane: `stream`, `scratch`, `pinned`, `pinEvent` are its
    void*        pinned[2]           = {};   // host staging blocks (tables built in place, one copy per launch)
    size_t       pinnedBytes[2]      = {};
    cudaEvent_t  pinEvent[2]         = {};   // recorded behind the copy out of a staging block: reservePinned waits
                                             // for it before the block is written again
    static constexpr int kGatherEvents = 9;  // mixture-sharded scoring (gmm_gather.cu): push stream, one event per
    cudaStream_t gatherStream        = nullptr;  // column range + one for the join, epoch of the flag words
    cudaEvent_t  gatherEvent[kGatherEvents] = {};
    struct GatherBinding {               // a gathered block this handle has called on: its flag words and the epoch of
        uint32_t* flags = nullptr;       // the last call that was fully enqueued (read from the block's own epoch word
        uint32_t  epoch = 0;             // when the handle first meets the block)
        // a call whose exchange has been started but not joined yet (gmm_score_device_gather_begin / gmm_gather_finish)
        bool      pending = false;
        void*     blocks[8] = {};
        uint32_t  rank = 0, world = 0, T = 0, col0 = 0, col1 = 0;
        size_t    ld = 0, flagsOffset = 0;
    };
    static constexpr int kGatherBindings = 4;
    GatherBinding gatherBound[kGatherBindings];
    int           gatherBoundNext  = 0;
    cudaEvent_t   gatherDone[kGatherBindings] = {};  // per binding: the last push of its pending call
    uint32_t*     gatherStatus     = nullptr;  // page-locked, device-mapped: epoch of the first exchange that timed out
    uint32_t*     gatherStatusDev  = nullptr;
    // gmm_score_host: device-to-host copies of finished chunks run on a stream of their own while the next chunk is
    // scored; gmm_score_async / gmm_wait: two lanes in turn, one event and token per lane
    static constexpr int kCopyEvents = 8;
    cudaStream_t copyStream          = nullptr;
    cudaEvent_t  copyEvent[kCopyEvents + 1] = {};
    static constexpr int kAsyncLanes = 2;        // lanes 1 and 2 (kMaxLanes - 1 at most)
    uint64_t     asyncSeq            = 0;        // tokens handed out so far
    uint64_t     asyncToken[kAsyncLanes] = {};   // token of the call last enqueued on async lane i (0: none)
    cudaEvent_t  asyncEvent[kAsyncLanes] = {};   // recorded behind that call's last copy
    HostPool*    pool                = nullptr;  // host threads of the table builders (hostPool())
    std::vector<std::vector<uint32_t>> workerWords;  // ... and one work space per thread
    struct EmissionStamps {                           // per thread: slot of an emission in the segment being listed,
        uint32_t              counter = 0;            // valid while stamp[e] == counter
        std::vector<uint32_t> slotOf, stamp;
    };
    std::vector<EmissionStamps> workerStamps;
    HostPool&    hostPool();
    int  pinnedCopied(int slot);
    int  useLane(int which);
    int  reservePinned(int slot, size_t bytes, void** out);
    std::vector<void*> owned;       // device allocations of the float model
    std::vector<void*> ownedQuant;  // device allocations of the quantized model

    void setError(const std::string& s) { error = s; }
    int  reserve(int slot, size_t bytes, void** out);
    int  devAlloc(void** p, size_t bytes, bool quant = false);
    int  upload(void** p, const void* src, size_t bytes, bool quant = false);
};

enum { PIN_ALIGN = 0, PIN_SUBSET = 1 };
enum { SCR_X = 0, SCR_SCORE = 1, SCR_BEST = 2, SCR_Q = 3, SCR_QX = 4, SCR_A = 5, SCR_ROW = 6, SCR_MISC = 7, SCR_BASE = 8, SCR_TILES = 9, SCR_VTILES = 10, SCR_TAB = 11, SCR_POST = 12, SCR_COUNT = 13 };

// gmm_align_emission_lists' results, handed to gmm_align_launch by a caller that already needed them
struct AlignLists {
    std::vector<uint32_t> emOff, emList, arcSlot, scratch;
};
// Tables built in a page-locked block (reservePinned) -> their device copy, by a kernel that reads the block over PCIe
// instead of a DMA copy: in the pipelined entry points a DMA copy of a few MB would queue behind the other lane's
// 50 MB feature upload on the host-to-device engine and hold the kernels back for a millisecond.
int gmm_copy_tables(gmm_handle* h, void* d_dst, const void* pinnedSrc, size_t bytes);
// gmm_direct.cu
int gmm_direct_score(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld);
int gmm_direct_posteriors(gmm_handle* h, uint32_t T, const float* d_x, float* d_post);
int gmm_direct_mixture_posteriors(gmm_handle* h, uint32_t T, const float* d_x, const uint32_t* d_mixture, float* d_post,
                                  size_t ldP, float* d_score, uint32_t maxK);
// gmm_quant.cu
int gmm_quant_launch(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best, float* d_score,
                     uint8_t* d_qx, size_t ld);
// gmm_tc.cu
int gmm_align_launch(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                     const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                     const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                     const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState, float* pathScore,
                     const uint32_t** d_pathEmission = nullptr, bool compact = false, uint32_t nEmissions = 0,
                     struct AlignDeferred* defer = nullptr, const struct AlignLists* lists = nullptr,
                     bool slotOutput = false);  // slotOutput: pathEmission = position in the segment's list (labels)
// d[r * ld + c] = fl(scale * d[r * ld + c]) for c < cols, on the handle's stream (the scaling wrapper of the reference)
int gmm_scale_block(gmm_handle* h, float* d, size_t rows, size_t cols, size_t ld, float scale);
// With `defer` gmm_align_launch returns once everything is enqueued; gmm_align_finish (same lane active) copies the
// outputs to the host arrays and waits.
struct AlignDeferred {
    bool           pending = false;
    const uint8_t *dPathEmission = nullptr, *dPathState = nullptr, *dPathScore = nullptr, *dOrdered = nullptr;
    uint32_t       nFrames = 0, nSeg = 0;
    uint32_t *     pathEmission = nullptr, *pathState = nullptr;
    float*         pathScore = nullptr;
    // Baum-Welch groups: posterior rows [wRows x wCols] (device pitch ldWdev, host pitch ldWhost floats), total scores
    const float*   dWeights = nullptr;
    float*         weights = nullptr;
    size_t         ldWdev = 0, ldWhost = 0;
    uint32_t       wRows = 0, wCols = 0;
    const double*  dTotal = nullptr;
    double*        total = nullptr;
};
// device copies of a Baum-Welch launch's segment tables, for the E-step over its posterior rows
struct BwDevice {
    const uint32_t *frameOff = nullptr, *emOff = nullptr, *emList = nullptr;
    const double*   total = nullptr;
};
int gmm_align_finish(gmm_handle* h, AlignDeferred& d);
// the distinct emissions of every segment's automaton in order of first use (the column order of a compact score
// block: gmm_tc_score_subset writes it, gmm_align_launch(compact) reads it)
bool gmm_align_emission_lists(gmm_handle* h, uint32_t nSeg, const uint32_t* stateOffsets, const uint32_t* arcOffsets,
                              const uint32_t* arcEmission, uint32_t nEmissions, struct AlignLists& lists);
// gmm_bw.cu
int gmm_bw_launch(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                  const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                  const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                  const float* finalWeight, float* d_weights, size_t ldW, double* totalScore, bool compact,
                  uint32_t nEmissions, const struct AlignLists* lists, bool wait, struct BwDevice* dev = nullptr);
// E-step over posterior rows: frame t of segment s is accumulated into mixture emList[emOff[s] + j] with weight
// d_rowWeights[t * ldW + j] for every j whose weight is positive (device arrays; frameOff / emOff [nSeg + 1])
int gmm_accum_rows_launch(gmm_handle* h, uint32_t nSeg, uint32_t T, const float* d_x, const uint32_t* d_frameOff,
                          const uint32_t* d_emOff, const uint32_t* d_emList, const float* d_rowWeights, size_t ldW);
int gmm_lda_launch(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_base, float* d_x);
int gmm_accum_launch(gmm_handle* h, uint32_t T, const float* d_x, const uint32_t* d_mixture, const double* d_weight,
                     uint32_t* d_best);
int gmm_qtc_pack(gmm_handle* h, const uint32_t* offsets, const uint8_t* qmean, const int32_t* c2);
int gmm_qtc_score(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best, float* d_score,
                  uint8_t* d_qx, size_t ld);
int gmm_tc_pack(gmm_handle* h, const uint32_t* offsets, const float* mean, const float* isr, const uint32_t* covIdx,
                const float* w2, const float* logNorm, bool force);
int gmm_tc_score_tiles(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, size_t ld, uint32_t tile0,
                       uint32_t tile1, bool prep, uint32_t prepFrames = 0, uint32_t frameTile0 = 0);
int gmm_gather_check(gmm_handle* h, const char* who);  // GMM_ERR_TIMEOUT after a timed-out exchange, until reset
int gmm_gather_reset_impl(gmm_handle* h);
int gmm_gather_finish_impl(gmm_handle* h, void* d_own_block, size_t flagsOffsetBytes);
int gmm_gather_score(gmm_handle* h, uint32_t T, const float* d_x, uint32_t rank, uint32_t world,
                     void* const* d_blocks, size_t ld, uint32_t colOffset, size_t flagsOffsetBytes, uint32_t chunks, bool join = true);
int gmm_tc_score(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld);
// scores of a per-segment subset of mixtures: column j of segment s's rows = mixture emList[emOff[s] + j]
// (host arrays); GMM_ERR_UNSUPPORTED when the model is not regular (the caller scores everything instead)
bool gmm_tc_subset_capable(const gmm_handle* h);
uint32_t gmm_tc_subset_ld(const gmm_handle* h, uint32_t maxListLength);
int gmm_tc_score_subset(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const uint32_t* emOff,
                        const uint32_t* emList, const float* d_x, float* d_scores, size_t ld);
