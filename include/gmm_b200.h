/* gmm_b200.h -- C-ABI of the B200-native GMM emission-scoring engine.
 *
 * This is the drop-in boundary: the only thing the RASR-side scorer class
 * (rasr/GpuFeatureScorer.cc, an Mm::AssigningFeatureScorer registered in Mm::FeatureScorerFactory)
 * binds to.  Plain C types, no RASR and no torch types; the library owns all device memory, the caller
 * owns all host buffers.  Every entry point returns GMM_OK (0) or a negative error code; the message
 * is available from gmm_last_error().  There is no CPU fallback: without a CUDA device gmm_create fails.
 *
 * Which reference interface each entry point replaces (paths relative to /root/reference/src/Mm):
 *
 *   gmm_upload_float    GaussDiagonalMaximumFeatureScorer::init             GaussDiagonalMaximumFeatureScorer.cc:64-86
 *                       (tables from MixtureFeatureScorerElement.cc:21-33, CovarianceFeatureScorerElement.cc:20-51)
 *   gmm_score_*  (max)  GaussDiagonalMaximumFeatureScorer::calculateScoreAndDensity + distance   .cc:116-218
 *   gmm_score_*  (sum)  GaussDiagonalSumFeatureScorer::calculateScoreAndDensity                  .cc:239-289
 *   gmm_posteriors      GaussDiagonalSumFeatureScorer::calculateDensityPosteriorProbabilities    .cc:291-298
 *   gmm_upload_quant    SimdGaussDiagonalMaximumFeatureScorer::init/buildMixtureTable  SimdFeatureScorer.cc:64-104
 *                       BatchIntFeatureScorer::init (with gmm_quant_final_score_rule)  BatchFeatureScorer.cc:337-376
 *   gmm_quant_* (qx)    SimdGaussDiagonalMaximumFeatureScorer::Context ctor            SimdFeatureScorer.cc:22-35
 *                       FeatureScorerIntelOptimization::multiplyAndQuantize            IntelOptimization.cc:50-67
 *   gmm_quant_* (q)     quantizedScore + FeatureScorerIntelOptimization::distance      SimdFeatureScorer.cc:158-176,
 *                       SSE2L2NormCodeGenerator (JIT SSE2)                             SSE2CodeGenerator.cc:324-374
 *   gmm_quant_* (score) calculateScoreAndDensity: f32(0.5 * q / scalingSquared)        SimdFeatureScorer.cc:135-145
 *   batching            BatchFeatureScorerBase ring buffer (caller side)               BatchFeatureScorer.cc:40-105
 *
 * Layout conventions: all matrices row-major; "slot" g in [0,G) enumerates (mixture, density) pairs in
 * mixture order, offsets[m]..offsets[m+1] are the slots of mixture m (tying already flattened, as
 * BatchFeatureScorerBase::initialize does, BatchFeatureScorer.cc:62-75).  Scores are -log p(x|m),
 * lower is better (FeatureScorer.hh:39-40).
 */
#ifndef GMM_B200_H
#define GMM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gmm_handle gmm_handle;

enum {
    GMM_OK            = 0,
    GMM_ERR_ARG       = -1, /* invalid argument / model not uploaded */
    GMM_ERR_CUDA      = -2, /* CUDA runtime error (message in gmm_last_error) */
    GMM_ERR_NO_DEVICE = -3, /* no usable sm_100 device */
    GMM_ERR_UNSUPPORTED = -4,
    GMM_ERR_TIMEOUT   = -5  /* a peer of a mixture-sharded exchange did not show up (gmm_gather_status / _reset) */
};

/* scoring mode of the float engine */
enum {
    GMM_MODE_MAX = 0, /* maximum approximation: 0.5 * min_k(...)      (diagonal-maximum) */
    GMM_MODE_SUM = 1  /* full sum: -log sum_k exp(-s_k)               (GaussDiagonalSumFeatureScorer) */
};

/* kernel selection for the float engine */
enum {
    GMM_KERNEL_AUTO   = 0, /* tensor-core kernel when the model fits it, else direct */
    GMM_KERNEL_DIRECT = 1, /* CUDA-core kernel, ((mu - x) * isr)^2 evaluated as the reference does */
    GMM_KERNEL_TENSOR = 2  /* tcgen05 split-fp16 contraction with fused min / log-sum-exp epilogue */
};

/* ---- lifetime ------------------------------------------------------------------------------- */
int         gmm_create(gmm_handle** out, int device);
void        gmm_destroy(gmm_handle* h);
const char* gmm_last_error(const gmm_handle* h); /* h may be NULL: error of the last failed gmm_create */
/* Run all work of this handle on an existing CUDA stream (a cudaStream_t passed as void*). NULL = own stream. */
int gmm_set_stream(gmm_handle* h, void* cuda_stream);

/* ---- float engine --------------------------------------------------------------------------- */
/* mean [G x D], isr [C x D] (inverse standard deviation, already scaled), cov_idx [G] in [0,C),
 * w2 [G] = minus2LogWeight of the slot (+inf for a zero-weight density), log_norm [C] = logNormFactor:
 * the four tables GaussDiagonalMaximumFeatureScorer::init builds, flattened per slot. */
int gmm_upload_float(gmm_handle* h, uint32_t M, uint32_t G, uint32_t D, uint32_t C, const uint32_t* offsets,
                     const float* mean, const float* isr, const uint32_t* cov_idx, const float* w2,
                     const float* log_norm, int mode, int kernel);

/* Score T frames held in HOST memory: x [T x D] -> scores [T x ld] (ld >= M), best [T x ld] or NULL.
 * Copies H2D, scores in chunks, copies each finished chunk D2H while the next one is scored, and waits.  best = index of the winning
 * density within its mixture (first minimum), 0xFFFFFFFF for a mixture without a finite density. */
int gmm_score_host(gmm_handle* h, uint32_t T, const float* x, float* scores, uint32_t* best, size_t ld);
/* The asynchronous pair (precedent: the buffered scorers score a batch while search consumes the
 * previous one, src/Nn/BatchFeatureScorer.cc:92-170): gmm_score_async enqueues what gmm_score_host does -- copy in,
 * score, copy out -- and returns at once with a token; gmm_wait(token) returns when that call's scores (and best) are
 * in the caller's arrays.  x, scores and best must stay valid until then, and should be page-locked (gmm_host_alloc):
 * with pageable memory the copies degrade to synchronous ones.  Two calls may be in flight (each on a lane of its own:
 * stream and device scratch); a third waits for the first.  Between a gmm_score_async and its gmm_wait the
 * handle takes gmm_score_async, gmm_wait, gmm_sync (which also waits for every outstanding token), and the blocking
 * gmm_score_host / gmm_mixture_posteriors_host, which work on the handle's own lane (the RASR plugin scores a few
 * late frames that way while a batch is in flight); no model upload and no other scoring entry point. */
int gmm_score_async(gmm_handle* h, uint32_t T, const float* x, float* scores, uint32_t* best, size_t ld,
                    uint64_t* token);
int gmm_wait(gmm_handle* h, uint64_t token);
/* Same with DEVICE pointers; asynchronous on the handle's stream (call gmm_sync to wait). */
int gmm_score_device(gmm_handle* h, uint32_t T, const float* d_x, float* d_scores, uint32_t* d_best, size_t ld);
/* Density posteriors of the sum scorer for T frames: post [T x G] (device pointers), exp(score - s_k). */
int gmm_posteriors_device(gmm_handle* h, uint32_t T, const float* d_x, float* d_post);

/* The same for ONE mixture per frame, host arrays: what AssigningContextScorer::getDensityPosteriorProbabilities(e, result)
 * (AssigningFeatureScorer.hh:46-47,126-128) asks of the sum scorer when AbstractMixtureSetEstimator accumulates in
 * baum-welch mode (AbstractMixtureSetEstimator.cc:134-146,386-395).  Frame t: post[t * ldP + k] = exp(score - s_k) for the
 * densities k of mixture[t] (0 beyond the mixture's densities; a frame with mixture 0xFFFFFFFF is skipped), and
 * score[t] (may be NULL) = the sum scorer's score of that mixture, evaluated in the reference's operation order.
 * ldP >= the largest mixture's density count. */
int gmm_mixture_posteriors_host(gmm_handle* h, uint32_t T, const float* x, const uint32_t* mixture, float* post,
                                size_t ldP, float* score);

/* ---- quantized engine (bit-exact SIMD-diagonal-maximum) -------------------------------------- */
/* qmean [G x Dpad] u8, cov_idx [G], qconst [G] s32, isr_scaled [C x D] f32: the tables exactly as the
 * reference's init() produced them on the host (so build-dependent rsqrt effects are inherited). */
int gmm_upload_quant(gmm_handle* h, uint32_t M, uint32_t G, uint32_t D, uint32_t Dpad, uint32_t C,
                     const uint32_t* offsets, const uint8_t* qmean, const uint32_t* cov_idx, const int32_t* qconst,
                     const float* isr_scaled, float scaling_squared);
/* How q becomes the final f32 score, per uploaded quantized model (gmm_upload_quant resets it to SIMD):
 *   GMM_QUANT_RULE_SIMD       f32(0.5 * q / s^2) evaluated in IEEE double   SimdFeatureScorer.cc:135-145
 *   GMM_QUANT_RULE_BATCH_INT  f32(q) / f32(2 s^2), one f32 division         BatchFeatureScorer.cc:468
 * The two agree for |q| < 2^24; they differ beyond (f32(q) rounds first), e.g. for the 0x80000000 constant of a
 * zero-weight density.  With the tables of BatchIntFeatureScorer::init (BatchFeatureScorer.cc:337-376; scaling_squared
 * = scale^2) the second rule makes the engine the reference's `batch-diagonal-maximum-int` scorer, bit for bit. */
#define GMM_QUANT_RULE_SIMD 0
#define GMM_QUANT_RULE_BATCH_INT 1
int gmm_quant_final_score_rule(gmm_handle* h, int rule);
/* x [T x D] f32 -> q [T x ld] s32 (may be NULL), best [T x ld] (may be NULL), score [T x ld] f32 (may
 * be NULL; f32(0.5 * q / scalingSquared) evaluated in IEEE double), qx [T x C x Dpad] u8 (may be NULL). */
int gmm_quant_score_host(gmm_handle* h, uint32_t T, const float* x, int32_t* q, uint32_t* best, float* score,
                         uint8_t* qx, size_t ld);
int gmm_quant_score_device(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best,
                           float* d_score, uint8_t* d_qx, size_t ld);

/* gmm_quant_score_device without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments and bit-identical results (q, best, score as gmm_quant_score_device; it always runs
 * the dp4a kernel), but the one scratch array the kernel needs -- the quantized frames, T * C * Dpad bytes -- lives in
 * a file-scope `static __device__` array of GMM_QUANT_STATIC_SCRATCH_BYTES that every GPU which loads the library
 * holds until exit, and the call is two kernel launches on the handle's stream: no cudaMalloc / cudaFree / allocating
 * library routine.  (The first call on a device also resolves the array's device address, cudaGetSymbolAddress:
 * make that call once before a stream capture begins.)
 * Limit: T * C * Dpad <= GMM_QUANT_STATIC_SCRATCH_BYTES (87 381 frames at C = 1, Dpad = 48); more is rejected with
 * GMM_ERR_ARG and a message naming the limit.  All calls on a device share the array and it keeps the last call's
 * bytes (every byte read is rewritten first, nothing needs clearing): calls whose kernels OVERLAP on the GPU --
 * different streams or handles of one device -- race; calls ordered on one stream do not. */
#define GMM_QUANT_STATIC_SCRATCH_BYTES (4u << 20)
int gmm_quant_score_device_static(gmm_handle* h, uint32_t T, const float* d_x, int32_t* d_q, uint32_t* d_best,
                                  float* d_score, size_t ld);

/* ---- mixture-sharded variant (BASELINE.json configs[4]): scoring with its exchange step --------------------------
 * Rank r of N holds mixtures [r*Mp, (r+1)*Mp) and scores the same frames as every other rank; afterwards every rank
 * needs every column.  The reference has no multi-GPU scorer (Mm::FeatureScorer::Scorer::score(e) is answered from
 * one process' own mixture set, src/Mm/FeatureScorer.hh:71-116); this is the all-gather configs[4] names, done by
 * the scoring call itself over peer-mapped memory (CUDA IPC; NVLink / NVSwitch), without NCCL or a host barrier.
 * d_blocks[0..world): every rank's block (own at index rank, the others mapped with gmm_peer_open): score rows
 * [maxFrames x ld] f32 followed, flagsOffsetBytes (a multiple of 128, >= T*ld*4) behind the start, by 32 flag words
 * (gmm_peer_alloc zeroes them).  The call (1) tells the peers that this rank's earlier work on the handle's stream
 * has finished reading its block, (2) scores the shard into columns [colOffset, colOffset + M) of the own block, one
 * range at a time (`chunks` ranges, 1..8: frame ranges when a range has at least 1024 frames, else ranges of B tiles
 * = columns, which need a model whose mixtures all have the same number of densities, else one range), and pushes
 * each finished range into the peers' blocks from a second stream while the next range is being scored, (3) makes the handle's stream wait until every peer's columns have landed.  Work
 * queued on the stream afterwards sees the complete [T x ld] rows; all ranks must make the same sequence of calls.
 * Tensor kernel only, T <= 32768.
 * Flag words: [r] READY from rank r, [8 + r] DONE from rank r, [16] CTA counter, [17] time-out marker, [18] the number
 * of calls this rank has made on the block.  The call counter lives with the block, not with the handle: a new
 * handle over used blocks (model re-upload, restarted rank) reads it at its first call and stays in step with its
 * peers; it advances only when a call was enqueued completely, so a call that returned an error is simply repeated.
 * A peer that does not show up within ~2 s (GMM_GATHER_TIMEOUT_MS) ends the wait.  This is neither silent nor
 * permanent: the waiting kernel (a) sets word 17 and stops announcing this rank's columns as complete, so the peers
 * time out too, (b) overwrites every column of the call's rows that is not this rank's own with +inf, so a consumer
 * already queued behind the call sees impossible scores, never stale ones, (c) stores the call's number into a
 * page-locked word: gmm_gather_status reads it (0 = no time-out; meaningful after a gmm_sync), and from then on
 * gmm_score_device_gather, gmm_align_device and gmm_baum_welch_device return GMM_ERR_TIMEOUT on this handle until
 * gmm_gather_reset.  (With one peer, long ranges are pushed by the copy engines behind a kernel that waits for the
 * peer's READY; if THAT wait times out the copies still run -- into the block of a peer that is already two seconds
 * late and will find no DONE from this rank -- and the condition is reported the same way.)  The reset is collective by contract: every rank calls it between two host barriers (no call in
 * flight anywhere); it zeroes the flag words of the blocks this handle has called on and restarts their numbering. */
int gmm_score_device_gather(gmm_handle* h, uint32_t T, const float* d_x, uint32_t rank, uint32_t world,
                            void* const* d_blocks, size_t ld, uint32_t colOffset, size_t flagsOffsetBytes,
                            uint32_t chunks);
/* The same call in two halves, for overlap ACROSS calls ("overlap tile i's gather with tile
 * i+1's GEMM"): _begin does steps (1) and (2) and returns with the last range's push still running on the second
 * stream; gmm_gather_finish(own block) is step (3).  With two block sets used in turn --
 *     begin(A, x0); begin(B, x1); finish(A) -> consume A; begin(A, x2); finish(B) -> consume B; ...
 * -- the last push of a call runs under the next call's scoring, and a call costs max(scoring, exchange) instead of
 * their sum.  A block with a started call is joined automatically when the next call on it begins. */
int gmm_score_device_gather_begin(gmm_handle* h, uint32_t T, const float* d_x, uint32_t rank, uint32_t world,
                                  void* const* d_blocks, size_t ld, uint32_t colOffset, size_t flagsOffsetBytes,
                                  uint32_t chunks);
int gmm_gather_finish(gmm_handle* h, void* d_own_block, size_t flagsOffsetBytes);
int gmm_gather_status(gmm_handle* h, uint32_t* timedOutCall);
int gmm_gather_reset(gmm_handle* h);
/* Device memory other processes of the same box can map: allocate (zeroed) + export (ipcHandle: 64 bytes), map,
 * unmap, free. */
int gmm_peer_alloc(gmm_handle* h, size_t bytes, void** d_ptr, void* ipcHandle);
int gmm_peer_open(gmm_handle* h, const void* ipcHandle, void** d_ptr);
int gmm_peer_close(gmm_handle* h, void* d_ptr);
int gmm_peer_free(gmm_handle* h, void* d_ptr);

/* ---- subset scoring ---------------------------------------------------------------------------------------------
 * The reference's context scorers evaluate a mixture when a consumer first asks for it (CachedContextScorer::score,
 * src/Mm/FeatureScorer.hh:170-230): an aligner pays for the few hundred emissions of the segment's automaton, not
 * for all M.  Same here: rows [frameOffsets[s], frameOffsets[s+1]) of d_scores receive, in column j, the score of
 * mixture mixtures[listOffsets[s] + j] -- bit-identical to column mixtures[..] of gmm_score_device on the tensor
 * kernel.  ld >= gmm_score_subset_ld(h, longest list); columns between a list's end and that bound are scratch.
 * frameOffsets / listOffsets / mixtures are HOST arrays.  GMM_ERR_UNSUPPORTED unless the tensor kernel is active
 * and all mixtures have the same number of densities (score everything then). */
int gmm_score_subset_device(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const uint32_t* listOffsets,
                            const uint32_t* mixtures, const float* d_x, float* d_scores, size_t ld);
uint32_t gmm_score_subset_ld(const gmm_handle* h, uint32_t maxListLength);

/* ---- device-resident consumer: Viterbi alignment over the score block ------------------
 * Replaces, for unpruned Viterbi alignment, Search::Aligner::SearchSpace::{addStartupHypothesis, expand,
 * activateOrUpdateStateHypothesisViterbi, getAlignmentFsaViterbi} (src/Search/Aligner.cc:132-158,190-239,398-437):
 * only the best path leaves the GPU (4 bytes per frame instead of 4*M).
 * nSeg independent segments; segment i owns rows [frameOffsets[i], frameOffsets[i+1]) of the score block and the
 * automaton states [stateOffsets[i], stateOffsets[i+1]).  Automata are given like Fsa states: arcs of (global)
 * state s are arcOffsets[s] .. arcOffsets[s+1], each with a SEGMENT-LOCAL target state, an emission index (column
 * of the score block; RASR: acousticModel->emissionIndex(arc.input)) and a transition weight; every arc consumes one
 * frame.  initialState[i] is segment-local; finalWeight[s] = +inf marks a non-final state.  All automaton and
 * output arrays are HOST memory.  Outputs: pathEmission[t] (emission index of frame t on the best path, 0xFFFFFFFF
 * for a segment that cannot reach a final state), pathState[t] (state entered at frame t; may be NULL),
 * pathScore[i] (best final score, FLT_MAX if none).  The call returns after the outputs are valid. */
int gmm_align_device(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                     const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                     const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                     const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState, float* pathScore);
/* Segments (since gmm_create) whose alignment met an exact f32 tie and was therefore replayed in the reference's
 * hypothesis order by the ordered pass (see gmm_align.cu); all other segments have a unique best path. */
uint64_t gmm_align_ordered_count(const gmm_handle* h);
/* RASR hands its aligner the scorers of the SCALED feature scorer: Mm::FeatureScorerScaling multiplies every score
 * by the acoustic model's scale before the aligner adds the transition weight (src/Mm/ScaledFeatureScorer.hh:62-64, one
 * f32 product).  With a scale != 1 (default 1) every gmm_score_*align*_host / gmm_score_baum_welch*_host call rounds
 * fl(scale * score) into its device-resident score block before the consumer reads it; gmm_score_host, gmm_score_device
 * and the *_device consumers (the caller's own block) are not affected. */
int gmm_set_consumer_score_scale(gmm_handle* h, float scale);
/* Features (host, [frameOffsets[nSeg] x D]) -> float scorer of this handle -> alignment, scores stay in HBM. */
int gmm_score_align_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                         const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                         const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                         const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState, float* pathScore);

/* The training pass: as gmm_score_align_host, then every aligned frame is accumulated into the estimators of
 * gmm_accumulator_create (below) with its path emission as mixture -- AlignmentNode -> KeyedEstimator in the
 * reference, here without the path leaving the device: pathEmission, pathState and pathScore may be NULL. */
int gmm_score_align_accumulate_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                                    const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                                    const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                                    const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState,
                                    float* pathScore);

/* ---- device-resident consumer, second mode: Baum-Welch alignment ------------------------------------------------
 * Replaces, for unpruned Baum-Welch alignment with min-weight 0 (the defaults), Search::Aligner in mode baum-welch:
 * SearchSpace::activateOrUpdateStateHypothesisBaumWelch / expand (src/Search/Aligner.cc:160-187, 190-239: forward
 * potentials collected in the log semiring), getAlignmentFsaBaumWelch + getAlignmentPosteriorFsa (:438-452, 757-772:
 * arc posteriors, Fsa::posterior64) and getAlignment (:693-716 with Speech/Alignment.cc:455-566: items combined per
 * frame and emission, clipped, exponentiated).  Automata as for gmm_align_device.
 * What comes back is, per frame, the weight of every alignment item the reference would emit: row t of `weights`
 * holds in column j the posterior of emission list[listOffsets[s] + j] of the frame's segment s (0 where the
 * reference has no item) -- the lists are the distinct emissions of each automaton in order of first use by an arc,
 * gmm_emission_lists computes them (host arrays; `list` needs room for one entry per arc).  ldW >= the longest list.
 * totalScore[s] (host, may be NULL) = -log p(segment | automaton), FLT_MAX and an all-zero block for a segment whose
 * automaton cannot reach a final state (the reference's empty alignment). */
int gmm_emission_lists(gmm_handle* h, uint32_t nSeg, const uint32_t* stateOffsets, const uint32_t* arcOffsets,
                       const uint32_t* arcEmission, uint32_t nEmissions, uint32_t* listOffsets, uint32_t* list);
/* scores and weights are DEVICE blocks ([frames x ld] as gmm_score_device wrote them, [frames x ldW]). */
int gmm_baum_welch_device(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_scores, size_t ld,
                          const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                          const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                          const float* finalWeight, float* d_weights, size_t ldW, double* totalScore);
/* Host features -> float scorer (only the emissions each automaton names) -> posteriors; weights is a HOST block. */
int gmm_score_baum_welch_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                              const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                              const uint32_t* arcEmission, const float* arcWeight, const uint32_t* initialState,
                              const float* finalWeight, float* weights, size_t ldW, double* totalScore);
/* The Baum-Welch training pass: as above, and every (frame, emission) item with a positive weight is accumulated
 * into the estimators of gmm_accumulator_create with that weight (AlignmentNode -> KeyedEstimator::processAlignedFeature
 * with weighted items, src/Speech/KeyedEstimator.cc:108-121); nothing returns per frame. */
int gmm_score_baum_welch_accumulate_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                                         const uint32_t* stateOffsets, const uint32_t* arcOffsets,
                                         const uint32_t* arcTarget, const uint32_t* arcEmission, const float* arcWeight,
                                         const uint32_t* initialState, const float* finalWeight, double* totalScore);

/* ---- the same two consumers for arcs that carry LABELS (what rasr/GpuAligner.cc binds) ----------------------------
 * The reference's alignment items name the arc's label -- an allophone state (Aligner.cc:205 / :185 store a->input_) --
 * and score it through acousticModel->emissionIndex(label) (Aligner.cc:225): several labels of one automaton may be
 * tied to ONE mixture, and the items keep them apart (Speech/Alignment.cc combineItems merges per (time, label)).
 * Here: arcLabel[a] is a SEGMENT-LOCAL label slot, labelMixture[labelOffsets[s] + slot] the mixture that scores it
 * (the same mixture may appear under several slots).  pathLabel[t] = slot of the arc taken at frame t (0xFFFFFFFF: no
 * final state reached); column j of row t of `weights` = posterior of label slot j.  Everything else as
 * gmm_score_align_host / gmm_score_baum_welch_host; the scores are bit for bit those of the emission-index calls
 * (a label's column is its mixture's column). */
int gmm_score_align_labels_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                                const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                                const uint32_t* arcLabel, const float* arcWeight, const uint32_t* initialState,
                                const float* finalWeight, const uint32_t* labelOffsets, const uint32_t* labelMixture,
                                uint32_t* pathLabel, uint32_t* pathState, float* pathScore);
int gmm_score_baum_welch_labels_host(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* x,
                                     const uint32_t* stateOffsets, const uint32_t* arcOffsets, const uint32_t* arcTarget,
                                     const uint32_t* arcLabel, const float* arcWeight, const uint32_t* initialState,
                                     const float* finalWeight, const uint32_t* labelOffsets,
                                     const uint32_t* labelMixture, float* weights, size_t ldW, double* totalScore);

/* ---- training E-step on the device ----------------------------------------------------
 * Replaces Speech::KeyedEstimator::processAlignedFeature (src/Speech/KeyedEstimator.cc:108-121) and
 * Mm::AbstractMixtureSetEstimator::accumulate (src/Mm/AbstractMixtureSetEstimator.cc:117-147): for every frame t
 * aligned to mixture[t] (0xFFFFFFFF: skip), the best density of that mixture under the uploaded float model's
 * scorer (reference operation order: the density index is the reference's bit for bit) and the accumulation of the
 * frame into that density's estimators (MixtureEstimator.cc:86-97, GaussDensityEstimator.cc:22-32,
 * VectorAccumulator.hh:60-67): mean sum += [w*]x, covariance sum += [w*]x*x, weights += w, all f64.
 * viterbi = 0 selects the posterior-weighted branch (sum scorer only): every density k with w * p_k > weightThreshold
 * (reference default: f32 epsilon) accumulates with weight w * p_k.
 * The estimators live on the device until gmm_accumulator_read; they belong to the float model (uploading a new
 * float model drops them).  meanIdx[g] names the mean of (mixture, density) slot g as the covariance table of
 * gmm_upload_float names its covariance; sums are [nMeans x D] / [C x D], mixWeight is [G]. */
int gmm_accumulator_create(gmm_handle* h, uint32_t nMeans, const uint32_t* meanIdx, int viterbi, double weightThreshold);
int gmm_accumulator_reset(gmm_handle* h);
int gmm_accumulate_device(gmm_handle* h, uint32_t T, const float* d_x, const uint32_t* d_mixture, const double* d_weight,
                          uint32_t* d_best);
int gmm_accumulate_host(gmm_handle* h, uint32_t T, const float* x, const uint32_t* mixture, const double* weight,
                        uint32_t* best);
/* The estimators as one block of gmm_accumulator_size() doubles, in the order meanSum [nMeans x D], meanWeight
 * [nMeans], covSum [C x D], covWeight [C], mixWeight [G]: copied to / from DEVICE memory on the handle's stream, so
 * that ranks can sum their estimators with one all-reduce (the combine step of
 * AbstractMixtureSetEstimator::accumulate(const AbstractMixtureSetEstimator&), .cc:149-169) without a host trip. */
size_t gmm_accumulator_size(const gmm_handle* h);
int gmm_accumulator_export_device(gmm_handle* h, double* d_dst);
int gmm_accumulator_import_device(gmm_handle* h, const double* d_src);
int gmm_accumulator_read(gmm_handle* h, double* meanSum, double* meanWeight, double* covSum, double* covWeight,
                         double* mixWeight);

/* ---- front end fused into the frame upload ---------------------------------------------
 * Replaces the reference's lda.flow network (src/Tools/FeatureExtraction/share/lda.flow): sliding window
 * (signal-vector-f32-sequence-concatenation: `length` frames, `right` of them after the current one, one output
 * per input frame, a missing neighbour = the closest frame of the segment; src/Signal/VectorSequenceConcatenation.hh:
 * 85-100, src/Signal/Delay.hh:32-46) followed by signal-matrix-multiplication-f32 (src/Signal/MatrixMult.hh:70;
 * out[n] = sum_i matrix[n][i] * in[i] in f32, left to right: src/Math/Matrix.hh:482-489, Math/Vector.hh:95-101;
 * bit-identical to the reference build).  matrix is [D x baseDim*length], row major, host memory.
 * Once set, every *_host entry point takes BASE features [T x baseDim] and transforms them on the device (windows
 * stay inside a segment; entry points without segments treat the call as one segment).  matrix = NULL clears. */
int gmm_set_lda(gmm_handle* h, uint32_t baseDim, uint32_t length, uint32_t right, uint32_t D, const float* matrix);
/* d_base [frames x baseDim] -> d_x [frames x D], both device; frameOffsets: host, nSeg + 1 entries. */
int gmm_lda_device(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_base, float* d_x);

/* ---- plumbing -------------------------------------------------------------------------------- */
/* Page-locked host buffers for the feature / score blocks (cudaMallocHost); NULL on failure. */
void* gmm_host_alloc(size_t bytes);
void  gmm_host_free(void* p);
int      gmm_sync(gmm_handle* h);                       /* wait for the handle's stream */
uint64_t gmm_launch_count(const gmm_handle* h);         /* kernels launched by this handle so far */
/* Measurement aid (bench.py's roofline): with timing on, every launch of a SCORING kernel (tensor, direct or
 * quantized scorer; not the frame preparation / quantization passes) is bracketed by CUDA events on the handle's
 * stream.  gmm_kernel_time waits for the stream, returns the summed device time and the number of bracketed
 * launches since the last call, and resets both. */
int gmm_kernel_timing(gmm_handle* h, int on);
int gmm_kernel_time(gmm_handle* h, double* total_ms, uint64_t* launches);
/* 1 = direct, 2 = tensor: the float kernel the last gmm_upload_float selected. */
int         gmm_float_kernel(const gmm_handle* h);
/* Tensor kernel only: MMA products per algorithmic product of the uploaded float model -- 2 (fp16 hi.hi + one e4m3
 * product for both corrections: the default for pooled models that pass the packing-time accuracy guard) or 3 (fp16
 * hi.hi + lo.hi + hi.lo: everything else, and always with GMM_TC_SPLIT=fp16 set at upload); 0 when the model is not on
 * the tensor kernel. */
int         gmm_float_split(const gmm_handle* h);
/* 1: the uploaded float model keeps the 1e-4 relative bar on the kernel it runs on.  0 only when the tensor kernel
 * was FORCED (kernel = GMM_KERNEL_TENSOR at upload) for a model whose centred, whitened means lie further out than
 * 8 sigma rms -- the automatic choice sends such a model to the direct kernel; a caller that forces the tensor
 * kernel (the mixture-sharded scorer has no direct variant) must check this and refuse or warn. */
int         gmm_float_accuracy_ok(const gmm_handle* h);
/* 1 = dp4a (CUDA cores), 2 = integer tensor path: the kernel the last gmm_upload_quant selected (0: no model).
 * Both are bit-identical to the reference; environment GMM_QUANT_KERNEL=dp4a at upload time forces the former. */
int         gmm_quant_kernel(const gmm_handle* h);
const char* gmm_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GMM_B200_H */
