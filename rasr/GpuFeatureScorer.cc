/** RASR-side plugin: see GpuFeatureScorer.hh.  Everything numeric happens behind the C-ABI; this file packs the
 *  Mm::MixtureSet with the reference's own element classes (so the tables are the reference's, including its
 *  build-dependent inverse square roots) and runs the buffered-scorer state machine.
 */
#include "GpuFeatureScorer.hh"

#include <cstring>

#include <Mm/CovarianceFeatureScorerElement.hh>
#include <Mm/IntelOptimization.hh>
#include <Mm/MixtureFeatureScorerElement.hh>
#include <Mm/Utilities.hh>

#include <gmm_b200.h>

using namespace MmGpu;

const Core::ParameterInt GpuFeatureScorer::paramBufferSize(
        "buffer-size", "frames scored per GPU batch in streaming mode; 0 = whole segments", 0, 0);
const Core::ParameterInt GpuFeatureScorer::paramDevice("device", "CUDA device index", 0, 0);
const Core::ParameterInt GpuFeatureScorer::paramKernel(
        "kernel", "float kernel: 0 = automatic, 1 = direct (reference operation order), 2 = tensor core", 0, 0, 2);
const Core::ParameterFloat GpuFeatureScorer::paramMixtureWeightScale(
        "mixture-weight-scale", "scaling of the logarithmized mixture weights", 1.0, Core::Type<Mm::Score>::epsilon);
const Core::ParameterFloat GpuFeatureScorer::paramGaussianScale(
        "gaussian-scale", "scaling of the logarithmized gaussian probability", 1.0, Core::Type<Mm::Score>::epsilon);
const Core::ParameterBool GpuFeatureScorer::paramDensityAssignment(
        "with-density-assignment",
        "compute the best density of every mixture with every batch (training consumers ask for score AND bestDensity, "
        "src/Mm/AbstractMixtureSetEstimator.cc:377-393); otherwise it costs a second pass when first asked for",
        false);

const Core::ParameterBool GpuFeatureScorer::paramDeferredScoring(
        "deferred-scoring",
        "send no batch to the GPU before a score of one of its frames is asked for (default: batches go out while the "
        "caller is still handing in frames).  For segments whose scorers go to MmGpu::GpuAligner, which scores on the "
        "device only what its automaton names: the full score matrix is then never computed",
        false);

// ---- score blocks ----------------------------------------------------------------------------------------

GpuFeatureScorer::PinnedPool::~PinnedPool() {
    for (size_t i = 0; i < free_.size(); ++i)
        gmm_host_free(free_[i].ptr);
}

void* GpuFeatureScorer::PinnedPool::acquire(size_t bytes, size_t& capacity) {
    size_t bestI = free_.size();
    for (size_t i = 0; i < free_.size(); ++i)  // smallest buffer that fits
        if (free_[i].bytes >= bytes && (bestI == free_.size() || free_[i].bytes < free_[bestI].bytes))
            bestI = i;
    if (bestI != free_.size()) {
        void* p  = free_[bestI].ptr;
        capacity = free_[bestI].bytes;
        free_.erase(free_.begin() + bestI);
        return p;
    }
    capacity = bytes + bytes / 4 + 4096;  // segments differ in length: leave room so that the next one fits too
    return gmm_host_alloc(capacity);
}

void GpuFeatureScorer::PinnedPool::release(void* ptr, size_t capacity) {
    if (!ptr)
        return;
    if (free_.size() >= 12) {  // bounded: a few batches in flight plus the blocks a caller holds on to
        gmm_host_free(ptr);
        return;
    }
    Buffer b = {ptr, capacity};
    free_.push_back(b);
}

GpuFeatureScorer::ScoreBlock::ScoreBlock(Core::Ref<PinnedPool> pl, u32 cap, u32 M, u32 D)
        : pool(pl), scores(0), best(0), features(0), scoresCap(0), bestCap(0), featuresCap(0), nFrames(0), capacity(cap),
          dimension(D),
          ld((M + 3) / 4 * 4),  // rows padded to 16 bytes: aligned row pieces on the device (include/gmm_b200.h)
          token(0) {
    features = static_cast<float*>(pool->acquire(size_t(capacity) * D * sizeof(float), featuresCap));
    capacity = featuresCap / (D * sizeof(float));
}

/** Next frame of the batch, written where the H2D copy will read it.  A full block grows by half (segment mode:
 *  the length of a segment is not known in advance). */
bool GpuFeatureScorer::ScoreBlock::append(const Mm::FeatureVector& f) {
    if (nFrames == capacity) {
        size_t cap   = 0;
        float* grown = static_cast<float*>(pool->acquire((size_t(capacity) + capacity / 2 + 64) * dimension * sizeof(float), cap));
        if (!grown)
            return false;
        memcpy(grown, features, size_t(nFrames) * dimension * sizeof(float));
        pool->release(features, featuresCap);
        features    = grown;
        featuresCap = cap;
        capacity    = cap / (dimension * sizeof(float));
    }
    if (!features)
        return false;
    std::copy(f.begin(), f.end(), features + size_t(nFrames) * dimension);
    ++nFrames;
    return true;
}

void GpuFeatureScorer::ScoreBlock::allocateScores() {
    if (!scores)
        scores = static_cast<Mm::Score*>(pool->acquire(size_t(nFrames) * ld * sizeof(Mm::Score), scoresCap));
}

void GpuFeatureScorer::ScoreBlock::allocateBest() {
    if (!best)
        best = static_cast<u32*>(pool->acquire(size_t(nFrames) * ld * sizeof(u32), bestCap));
}

GpuFeatureScorer::ScoreBlock::~ScoreBlock() {
    pool->release(scores, scoresCap);
    pool->release(best, bestCap);
    pool->release(features, featuresCap);
}

// ---- context scorer --------------------------------------------------------------------------------------

Mm::EmissionIndex GpuFeatureScorer::Context::nEmissions() const {
    return parent_->nMixtures();
}

/** The first request for a frame without scores sends its whole batch to the GPU; a batch that is still in flight
 *  (gmm_score_async) is waited for. */
inline void GpuFeatureScorer::Context::ready(bool withBest) const {
    if (!block_->scores)
        parent_->scoreBlock(*block_, withBest);
    if (block_->token)
        parent_->complete(*block_);
}

Mm::Score GpuFeatureScorer::Context::score(Mm::EmissionIndex e) const {
    require_(0 <= e && e < nEmissions());
    ready(false);
    return block_->scores[size_t(row_) * block_->ld + e];
}

Mm::DensityInMixture GpuFeatureScorer::Context::bestDensity(Mm::EmissionIndex e) const {
    require_(0 <= e && e < nEmissions());
    ready(true);
    if (!block_->best)
        parent_->assignDensities(*block_);
    return block_->best[size_t(row_) * block_->ld + e];
}

/** Not offered by the reference's diagonal scorers either: CachedAssigningFeatureScorer::calculateScore(cs, e, dns) is
 *  a criticalError unless a subclass overrides it (src/Mm/AssigningFeatureScorer.hh:131-134), and neither
 *  GaussDiagonalMaximumFeatureScorer nor GaussDiagonalSumFeatureScorer does. */
Mm::Score GpuFeatureScorer::Context::score(Mm::EmissionIndex, Mm::DensityIndex) const {
    parent_->criticalError("This feature scorer does not support the calculation of scores given a density in the mixture");
    return Core::Type<Mm::Score>::max;
}

/** The sum scorer's density posteriors exp(score - s_k) (GaussDiagonalMaximumFeatureScorer.cc:291-298), which
 *  AbstractMixtureSetEstimator accumulates in baum-welch mode (AbstractMixtureSetEstimator.cc:134-146,386-395).  The
 *  maximum and the quantized scorers answer with the reference's criticalError (AssigningFeatureScorer.hh:135-138). */
void GpuFeatureScorer::Context::getDensityPosteriorProbabilities(Mm::EmissionIndex e, std::vector<Mm::Weight>& result) const {
    require_(0 <= e && e < nEmissions());
    if (parent_->kind_ != kindSum) {
        parent_->criticalError("This feature scorer does not support the calculation of density posterior probabilities");
        return;
    }
    ready(false);
    parent_->densityPosteriors(*block_, row_, e, result);
}

// ---- construction: pack the mixture set --------------------------------------------------------------------

GpuFeatureScorer::GpuFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> mixtureSet, Kind kind,
                                   u32 restrictions)
        : Core::Component(c),
          Precursor(c),
          kind_(kind),
          handle_(0),
          bufferSize_(paramBufferSize(c)),
          withBest_(paramDensityAssignment(c)),
          deferred_(paramDeferredScoring(c)),
          pool_(new PinnedPool),
          nMixtures_(mixtureSet->nMixtures()),
          dimension_(mixtureSet->dimension()),
          nDensities_(0),
          nModelDensities_(mixtureSet->nDensities()),
          maxDensities_(0),
          nBatches_(0),
          nAssignmentPasses_(0) {
    if ((restrictions & pooledCovarianceOnly) && mixtureSet->nCovariances() != 1)
        criticalError("feature scorer supports only globally pooled covariance");  // as BatchFeatureScorer.cc:146-148
    if ((restrictions & atMost48PaddedComponents) && ((dimension_ + 15) / 16) * 16 > 48)
        criticalError("This feature scorer supports only features with max. %d components", 48);  // as :550-553
    if (gmm_create(&handle_, paramDevice(config)) != GMM_OK)
        criticalError("cannot create the GPU scoring engine: %s", gmm_last_error(0));
    // flat (mixture, density) slots, as BatchFeatureScorerBase::initialize (BatchFeatureScorer.cc:62-75)
    offsets_.resize(nMixtures_ + 1, 0);
    for (Mm::MixtureIndex m = 0; m < nMixtures_; ++m) {
        offsets_[m] = nDensities_;
        nDensities_ += mixtureSet->mixture(m)->nDensities();
        maxDensities_ = std::max<u32>(maxDensities_, mixtureSet->mixture(m)->nDensities());
    }
    offsets_[nMixtures_] = nDensities_;
    // mixture elements as GaussDiagonalMaximumFeatureScorer::init keeps them (.cc:72-79): weights scaled for the float
    // scorers, as they are for the quantized ones
    mixtureTable_.resize(nMixtures_);
    for (Mm::MixtureIndex m = 0; m < nMixtures_; ++m) {
        mixtureTable_[m] = *mixtureSet->mixture(m);
        if (!quantized())
            mixtureTable_[m].scale(paramMixtureWeightScale(config));
    }
    if (kind_ == kindQuantized)
        initQuantized(*mixtureSet);
    else if (kind_ == kindBatchInt)
        initBatchInt(*mixtureSet);
    else
        initFloat(*mixtureSet);
    log("GPU feature scorer: %d mixtures, %d densities, dimension %d, %s, buffer size %d", int(nMixtures_),
        int(nDensities_), int(dimension_),
        quantized() ? "u8 quantized" : (gmm_float_kernel(handle_) == GMM_KERNEL_TENSOR ? "tensor-core kernel" : "direct kernel"),
        int(bufferSize_));
}

GpuFeatureScorer::~GpuFeatureScorer() {
    fifo_.clear();
    staging_.reset();
    segment_.clear();
    gmm_destroy(handle_);
}

void GpuFeatureScorer::check(int rc, const char* what) const {
    if (rc != GMM_OK)
        criticalError("%s failed (%d): %s", what, rc, gmm_last_error(handle_));
}

/** tables of GaussDiagonalMaximumFeatureScorer::init (GaussDiagonalMaximumFeatureScorer.cc:64-86) */
void GpuFeatureScorer::initFloat(const Mm::MixtureSet& ms) {
    const Mm::Score        mixtureWeightScale = paramMixtureWeightScale(config);
    const Mm::VarianceType gaussianScale      = std::sqrt(paramGaussianScale(config));
    const size_t           D = dimension_, C = ms.nCovariances();
    std::vector<float>     isr(C * D), logNorm(C), mean(size_t(nDensities_) * D), w2(nDensities_);
    std::vector<u32>       covIdx(nDensities_);
    for (size_t c = 0; c < C; ++c) {
        Mm::CovarianceFeatureScorerElement e;
        e = *ms.covariance(c);
        e.scale(gaussianScale);
        std::copy(e.inverseSquareRootDiagonal().begin(), e.inverseSquareRootDiagonal().end(), isr.begin() + c * D);
        logNorm[c] = e.logNormalizationFactor();
    }
    for (Mm::MixtureIndex m = 0; m < nMixtures_; ++m) {
        Mm::MixtureFeatureScorerElement e;
        e = *ms.mixture(m);
        e.scale(mixtureWeightScale);
        for (size_t k = 0; k < e.nDensities(); ++k) {
            const size_t            g   = offsets_[m] + k;
            const Mm::GaussDensity* dns = ms.density(e.densityIndex(k));
            const Mm::Mean&         mu  = *ms.mean(dns->meanIndex());
            std::copy(mu.begin(), mu.end(), mean.begin() + g * D);
            covIdx[g] = dns->covarianceIndex();
            w2[g]     = e.minus2LogWeights()[k];
        }
    }
    check(gmm_upload_float(handle_, nMixtures_, nDensities_, D, C, &offsets_[0], &mean[0], &isr[0], &covIdx[0], &w2[0],
                           &logNorm[0], kind_ == kindSum ? GMM_MODE_SUM : GMM_MODE_MAX, paramKernel(config)),
          "gmm_upload_float");
}

/** tables of SimdGaussDiagonalMaximumFeatureScorer::init / getScaling / buildMixtureTable
 *  (SimdFeatureScorer.cc:64-133), built with the reference's own element classes */
void GpuFeatureScorer::initQuantized(const Mm::MixtureSet& ms) {
    typedef Mm::FeatureScorerIntelOptimization Opt;
    const size_t                                    D = dimension_, C = ms.nCovariances();
    std::vector<Mm::CovarianceFeatureScorerElement> cov(C);
    for (size_t c = 0; c < C; ++c)
        cov[c] = *ms.covariance(c);
    Mm::MeanType minMean = Core::Type<Mm::MeanType>::max, maxMean = Core::Type<Mm::MeanType>::min;
    for (Mm::DensityIndex i = 0; i < ms.nDensities(); ++i) {
        const Mm::GaussDensity*              dns = ms.density(i);
        const Mm::Mean&                      mu  = *ms.mean(dns->meanIndex());
        const std::vector<Mm::VarianceType>& is  = cov[dns->covarianceIndex()].inverseSquareRootDiagonal();
        for (Mm::ComponentIndex d = 0; d < mu.size(); ++d) {
            Mm::MeanType v = mu[d] * is[d];
            minMean        = std::min(minMean, v);
            maxMean        = std::max(maxMean, v);
        }
    }
    // quantizationScalingFactor (SimdFeatureScorer.cc:128-133)
    const int       quantizedIntervalSize = (int)Core::Type<Opt::QuantizedType>::max - (int)Core::Type<Opt::QuantizedType>::min;
    Mm::FeatureType intervalSize          = 2 * std::max(Core::abs(minMean), Core::abs(maxMean));
    Mm::MeanType    scaling               = (Mm::FeatureType)quantizedIntervalSize / (1.25 * intervalSize);
    const Mm::Score scalingSquared        = scaling * scaling;
    log("Scaling factor: ") << scaling;
    for (size_t c = 0; c < C; ++c)
        cov[c].scale(scaling);
    std::vector<float> isrS(C * D);
    for (size_t c = 0; c < C; ++c)
        std::copy(cov[c].inverseSquareRootDiagonal().begin(), cov[c].inverseSquareRootDiagonal().end(),
                  isrS.begin() + c * D);
    size_t           Dpad = 0;
    std::vector<u8>  qmean;
    std::vector<s32> qconst(nDensities_);
    std::vector<u32> covIdx(nDensities_);
    for (Mm::MixtureIndex m = 0; m < nMixtures_; ++m) {
        const Mm::Mixture* mixture = ms.mixture(m);
        for (size_t k = 0; k < mixture->nDensities(); ++k) {
            const size_t                 g   = offsets_[m] + k;
            const Mm::GaussDensity*      dns = ms.density(mixture->densityIndex(k));
            Opt::MixtureElement::Density q;
            Mm::Weight scaledMinus2LogWeight = scalingSquared * -2 * mixture->logWeight(k);
            Opt::createDensityElement(scaledMinus2LogWeight, *ms.mean(dns->meanIndex()), cov[dns->covarianceIndex()], q);
            if (qmean.empty()) {
                Dpad = q.preparedMean_.size();
                qmean.resize(size_t(nDensities_) * Dpad);
            }
            std::copy(q.preparedMean_.begin(), q.preparedMean_.end(), qmean.begin() + g * Dpad);
            qconst[g] = q.constantWeight_;
            covIdx[g] = dns->covarianceIndex();
        }
    }
    if (Dpad == 0)
        Dpad = ((D + 15) / 16) * 16;
    check(gmm_upload_quant(handle_, nMixtures_, nDensities_, D, Dpad, C, &offsets_[0], qmean.empty() ? 0 : &qmean[0],
                           &covIdx[0], &qconst[0], &isrS[0], scalingSquared),
          "gmm_upload_quant");
}

/** tables of BatchIntFeatureScorer::quantizationScale / init (BatchFeatureScorer.cc:319-376), from the same
 *  expressions under the same flags: the u8 grid of the SIMD scorer, constants (s32)(L s^2 - 2 s^2 logWeight), final
 *  score q / (2 s^2) (:468).  The device kernel is the quantized scorer's. */
void GpuFeatureScorer::initBatchInt(const Mm::MixtureSet& ms) {
    if (ms.nCovariances() != 1)
        criticalError("feature scorer supports only globally pooled variance");  // as BatchFeatureScorer.cc:340-342
    const size_t                       D = dimension_, Dpad = ((D + 15) / 16) * 16;
    Mm::CovarianceFeatureScorerElement covariance;
    covariance = *ms.covariance(0);
    std::vector<f32> variance(covariance.inverseSquareRootDiagonal().begin(), covariance.inverseSquareRootDiagonal().end());
    Mm::MeanType     minMean = Core::Type<Mm::MeanType>::max, maxMean = Core::Type<Mm::MeanType>::min;
    for (Mm::DensityIndex i = 0; i < ms.nDensities(); ++i) {
        const Mm::Mean& mu = *ms.mean(ms.density(i)->meanIndex());
        for (Mm::ComponentIndex d = 0; d < mu.size(); ++d) {
            Mm::MeanType v = mu[d] * variance[d];
            minMean        = std::min(minMean, v);
            maxMean        = std::max(maxMean, v);
        }
    }
    const int quantizedIntervalSize = (int)Core::Type<u8>::max - (int)Core::Type<u8>::min;
    f32       intervalSize          = 2 * std::max(Core::abs(minMean), Core::abs(maxMean));
    f32       scale                 = static_cast<f32>(quantizedIntervalSize) / (1.25 * intervalSize);
    f32       scaleSquared          = scale * scale;
    f32       scale2                = 2.0 * scaleSquared;
    for (size_t d = 0; d < D; ++d)
        variance[d] = variance[d] * scale;
    const float      logNormFactor = covariance.logNormalizationFactor() * scaleSquared;
    std::vector<u8>  qmean(size_t(nDensities_) * Dpad, 0);
    std::vector<s32> qconst(nDensities_);
    std::vector<u32> covIdx(nDensities_, 0);
    Mm::quantize<f32, u8> quantize;
    for (Mm::MixtureIndex m = 0; m < nMixtures_; ++m) {
        const Mm::Mixture& mixture = *ms.mixture(m);
        for (size_t k = 0; k < mixture.nDensities(); ++k) {
            const size_t    g  = offsets_[m] + k;
            const Mm::Mean& mu = *ms.mean(ms.density(mixture.densityIndex(k))->meanIndex());
            for (size_t d = 0; d < D; ++d)
                qmean[g * Dpad + d] = quantize(mu[d] * variance[d]);
            qconst[g] = static_cast<s32>(logNormFactor - scale2 * mixture.logWeight(k));
        }
    }
    check(gmm_upload_quant(handle_, nMixtures_, nDensities_, D, Dpad, 1, &offsets_[0], qmean.empty() ? 0 : &qmean[0],
                           &covIdx[0], &qconst[0], &variance[0], scale2 * 0.5f),
          "gmm_upload_quant");
    check(gmm_quant_final_score_rule(handle_, GMM_QUANT_RULE_BATCH_INT), "gmm_quant_final_score_rule");
}

// ---- buffered protocol (FeatureScorer.hh:86-134; state machine of BatchFeatureScorer.cc:40-105) -------------

void GpuFeatureScorer::reset() const {
    fifo_.clear();
    staging_.reset();  // contexts a caller still holds keep their block and can still be scored
    segment_.clear();
}

void GpuFeatureScorer::push(const Mm::FeatureVector& f) const {
    require(f.size() == dimension_);  // as GaussDiagonalMaximumFeatureScorer.cc:32
    if (!staging_)
        staging_ = Core::Ref<ScoreBlock>(new ScoreBlock(pool_, bufferSize_ ? bufferSize_ : 1024, nMixtures_, dimension_));
    if (!staging_->append(f))
        criticalError("cannot allocate page-locked memory for %d frames", int(staging_->nFrames + 1));
    ContextRef c(new Context(this, staging_, staging_->nFrames - 1));
    fifo_.push_back(c);
    if (bufferSize_ == 0)
        segment_.push_back(c);
    // Frames go to the GPU while the caller is still handing in (or consuming) others (gmm_score_async; the precedent
    // is src/Nn/BatchFeatureScorer.cc:92-170): in streaming mode half a buffer at a time, in segment mode every
    // kSegmentBatch frames, so that the scores of a long segment are already on their way over the host link when the
    // first one is asked for.  The float engine only.
    if (!quantized() && !deferred_ &&
        staging_->nFrames >= (bufferSize_ ? std::max<u32>(2, bufferSize_ / 2) : u32(kSegmentBatch)))
        scoreBlock(*staging_, withBest_, true);
}

void GpuFeatureScorer::addFeature(const Mm::FeatureVector& f) const {
    require(!bufferFilled());
    push(f);
}

Mm::AssigningFeatureScorer::AssigningScorer GpuFeatureScorer::getAssigningScorer(const Mm::FeatureVector& f) const {
    push(f);
    ContextRef c = fifo_.front();
    fifo_.pop_front();
    return AssigningScorer(c.get());
}

Mm::FeatureScorer::Scorer GpuFeatureScorer::flush() const {
    require(!bufferEmpty());
    // the caller has begun to drain the buffer (end of the segment): the frames handed in since the last batch go to
    // the GPU now, so that they are on their way while the earlier frames' scores are being read
    if (staging_ && staging_->nFrames > 0 && !quantized() && !deferred_)
        scoreBlock(*staging_, withBest_, true);
    ContextRef c = fifo_.front();
    fifo_.pop_front();
    return Scorer(c.get());
}

Mm::FeatureScorer::Scorer GpuFeatureScorer::getTimeIndexedScorer(u32 time) const {
    require(hasTimeIndexedCache());
    require(time < segment_.size());
    return Scorer(segment_[time].get());
}

/** One GPU batch over the frames of a block.  async: enqueue only (gmm_score_async); the first score() of one of its
 *  frames waits for the token.  The block stops taking frames: the next frame opens a new one. */
void GpuFeatureScorer::scoreBlock(ScoreBlock& block, bool withBest, bool async) const {
    if (staging_.get() == &block)
        staging_.reset();  // the contexts hold the block
    withBest = withBest || withBest_;
    block.allocateScores();
    if (withBest)
        block.allocateBest();
    if (!block.scores || (withBest && !block.best))
        criticalError("cannot allocate page-locked memory for %d x %d scores", int(block.nFrames), int(nMixtures_));
    const u32 T = block.nFrames;
    if (quantized())
        check(gmm_quant_score_host(handle_, T, block.features, 0, block.best, block.scores, 0, block.ld), "gmm_quant_score_host");
    else if (async) {
        uint64_t token = 0;
        check(gmm_score_async(handle_, T, block.features, block.scores, block.best, block.ld, &token), "gmm_score_async");
        block.token = token;
    }
    else
        check(gmm_score_host(handle_, T, block.features, block.scores, block.best, block.ld), "gmm_score_host");
    ++nBatches_;
}

/** The block's asynchronous call has finished (its scores are in host memory). */
void GpuFeatureScorer::complete(ScoreBlock& block) const {
    check(gmm_wait(handle_, block.token), "gmm_wait");
    block.token = 0;
}

/** bestDensity() on a block that was scored without the assignment: one more pass over its frames. */
void GpuFeatureScorer::assignDensities(ScoreBlock& block) const {
    block.allocateBest();
    if (!block.best)
        criticalError("cannot allocate page-locked memory for the density assignment");
    check(gmm_sync(handle_), "gmm_sync");  // no asynchronous call in flight while a blocking one runs
    ++nAssignmentPasses_;
    if (quantized())
        check(gmm_quant_score_host(handle_, block.nFrames, block.features, 0, block.best, 0, 0, block.ld),
              "gmm_quant_score_host");
    else
        check(gmm_score_host(handle_, block.nFrames, block.features, block.scores, block.best, block.ld), "gmm_score_host");
}

/** Density posteriors of emission e for one frame of a scored block (sum scorer). */
void GpuFeatureScorer::densityPosteriors(const ScoreBlock& block, u32 row, Mm::EmissionIndex e,
                                         std::vector<Mm::Weight>& result) const {
    const u32          K = offsets_[e + 1] - offsets_[e];
    std::vector<float> post(std::max<u32>(1, maxDensities_), 0.f);
    const uint32_t     mixture = e;
    check(gmm_sync(handle_), "gmm_sync");
    check(gmm_mixture_posteriors_host(handle_, 1, block.features + size_t(row) * dimension_, &mixture, &post[0],
                                      post.size(), 0),
          "gmm_mixture_posteriors_host");
    result.assign(post.begin(), post.begin() + K);
}
