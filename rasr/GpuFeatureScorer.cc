/** RASR-side plugin: see GpuFeatureScorer.hh.  Everything numeric happens behind the C-ABI; this file packs the
 *  Mm::MixtureSet with the reference's own element classes (so the tables are the reference's, including its
 *  build-dependent inverse square roots) and runs the buffered-scorer state machine.
 */
#include "GpuFeatureScorer.hh"

#include <cstring>

#include <Mm/CovarianceFeatureScorerElement.hh>
#include <Mm/IntelOptimization.hh>
#include <Mm/MixtureFeatureScorerElement.hh>

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

// ---- score blocks ----------------------------------------------------------------------------------------

GpuFeatureScorer::ScoreBlock::ScoreBlock(u32 T, u32 M)
        : scores(0), best(0), nFrames(T), nMixtures(M) {
    scores = static_cast<Mm::Score*>(gmm_host_alloc(size_t(T) * M * sizeof(Mm::Score)));
}

GpuFeatureScorer::ScoreBlock::~ScoreBlock() {
    gmm_host_free(scores);
    gmm_host_free(best);
}

// ---- context scorer --------------------------------------------------------------------------------------

GpuFeatureScorer::Context::Context(const GpuFeatureScorer* parent, const Mm::FeatureVector& f)
        : parent_(parent), row_(0), pending_(f.begin(), f.end()) {
    require(f.size() == parent->dimension());  // as GaussDiagonalMaximumFeatureScorer.cc:32
}

Mm::EmissionIndex GpuFeatureScorer::Context::nEmissions() const {
    return parent_->nMixtures();
}

Mm::Score GpuFeatureScorer::Context::score(Mm::EmissionIndex e) const {
    require_(0 <= e && e < nEmissions());
    if (!block_)
        parent_->scorePending(false);
    return block_->scores[size_t(row_) * block_->nMixtures + e];
}

Mm::DensityInMixture GpuFeatureScorer::Context::bestDensity(Mm::EmissionIndex e) const {
    require_(0 <= e && e < nEmissions());
    if (!block_)
        parent_->scorePending(true);
    if (!block_->best)
        parent_->assignDensities(*block_);
    return block_->best[size_t(row_) * block_->nMixtures + e];
}

Mm::Score GpuFeatureScorer::Context::score(Mm::EmissionIndex, Mm::DensityIndex) const {
    parent_->criticalError("score(emission, density) is not available from the GPU feature scorer");
    return Core::Type<Mm::Score>::max;
}

void GpuFeatureScorer::Context::getDensityPosteriorProbabilities(Mm::EmissionIndex, std::vector<Mm::Weight>&) const {
    parent_->criticalError("getDensityPosteriorProbabilities() is not available from the GPU feature scorer");
}

// ---- construction: pack the mixture set --------------------------------------------------------------------

GpuFeatureScorer::GpuFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> mixtureSet, Kind kind)
        : Core::Component(c),
          Precursor(c),
          kind_(kind),
          handle_(0),
          bufferSize_(paramBufferSize(c)),
          nMixtures_(mixtureSet->nMixtures()),
          dimension_(mixtureSet->dimension()),
          nDensities_(0),
          nBatches_(0) {
    if (gmm_create(&handle_, paramDevice(config)) != GMM_OK)
        criticalError("cannot create the GPU scoring engine: %s", gmm_last_error(0));
    // flat (mixture, density) slots, as BatchFeatureScorerBase::initialize (BatchFeatureScorer.cc:62-75)
    offsets_.resize(nMixtures_ + 1, 0);
    for (Mm::MixtureIndex m = 0; m < nMixtures_; ++m) {
        offsets_[m] = nDensities_;
        nDensities_ += mixtureSet->mixture(m)->nDensities();
    }
    offsets_[nMixtures_] = nDensities_;
    if (kind_ == kindQuantized)
        initQuantized(*mixtureSet);
    else
        initFloat(*mixtureSet);
    log("GPU feature scorer: %d mixtures, %d densities, dimension %d, %s, buffer size %d", int(nMixtures_),
        int(nDensities_), int(dimension_),
        kind_ == kindQuantized ? "u8 quantized" : (gmm_float_kernel(handle_) == GMM_KERNEL_TENSOR ? "tensor-core kernel" : "direct kernel"),
        int(bufferSize_));
}

GpuFeatureScorer::~GpuFeatureScorer() {
    fifo_.clear();
    pending_.clear();
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

// ---- buffered protocol (FeatureScorer.hh:86-134; state machine of BatchFeatureScorer.cc:40-105) -------------

void GpuFeatureScorer::reset() const {
    fifo_.clear();
    pending_.clear();
    segment_.clear();
}

void GpuFeatureScorer::push(const Mm::FeatureVector& f) const {
    ContextRef c(new Context(this, f));
    fifo_.push_back(c);
    pending_.push_back(c);
    if (bufferSize_ == 0)
        segment_.push_back(c);
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
    ContextRef c = fifo_.front();
    fifo_.pop_front();
    return Scorer(c.get());
}

Mm::FeatureScorer::Scorer GpuFeatureScorer::getTimeIndexedScorer(u32 time) const {
    require(hasTimeIndexedCache());
    require(time < segment_.size());
    return Scorer(segment_[time].get());
}

/** One GPU batch over every frame that has no scores yet. */
void GpuFeatureScorer::scorePending(bool withBest) const {
    const u32 T = pending_.size();
    if (T == 0)
        return;
    Core::Ref<ScoreBlock> block(new ScoreBlock(T, nMixtures_));
    if (!block->scores)
        criticalError("cannot allocate page-locked memory for %d x %d scores", int(T), int(nMixtures_));
    block->features.resize(size_t(T) * dimension_);
    for (u32 t = 0; t < T; ++t)
        std::copy(pending_[t]->pending_.begin(), pending_[t]->pending_.end(),
                  block->features.begin() + size_t(t) * dimension_);
    if (withBest) {
        block->best = static_cast<u32*>(gmm_host_alloc(size_t(T) * nMixtures_ * sizeof(u32)));
        if (!block->best)
            criticalError("cannot allocate page-locked memory for the density assignment");
    }
    if (kind_ == kindQuantized)
        check(gmm_quant_score_host(handle_, T, &block->features[0], 0, block->best, block->scores, 0, nMixtures_),
              "gmm_quant_score_host");
    else
        check(gmm_score_host(handle_, T, &block->features[0], block->scores, block->best, nMixtures_), "gmm_score_host");
    for (u32 t = 0; t < T; ++t) {
        pending_[t]->block_ = block;
        pending_[t]->row_   = t;
        std::vector<float>().swap(pending_[t]->pending_);
    }
    pending_.clear();
    ++nBatches_;
}

/** bestDensity() on a block that was scored without the assignment: one more pass over its frames. */
void GpuFeatureScorer::assignDensities(ScoreBlock& block) const {
    const size_t n = size_t(block.nFrames) * block.nMixtures;
    block.best     = static_cast<u32*>(gmm_host_alloc(n * sizeof(u32)));
    if (!block.best)
        criticalError("cannot allocate page-locked memory for the density assignment");
    if (kind_ == kindQuantized)
        check(gmm_quant_score_host(handle_, block.nFrames, &block.features[0], 0, block.best, 0, 0, block.nMixtures),
              "gmm_quant_score_host");
    else
        check(gmm_score_host(handle_, block.nFrames, &block.features[0], block.scores, block.best, block.nMixtures),
              "gmm_score_host");
}
