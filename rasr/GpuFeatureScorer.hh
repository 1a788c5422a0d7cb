/** RASR-side plugin: GPU GMM feature scorers behind Mm::FeatureScorerFactory.
 *
 *  Compiled INTO RASR (module MODULE_MM_GPU, see rasr/Module.hh and INTEGRATION.md); binds only to the C-ABI of
 *  include/gmm_b200.h (libgmm_b200.so, hand-written CUDA for sm_100a).  Six names, drop-in for
 *
 *    gpu-diagonal-maximum        Mm::GaussDiagonalMaximumFeatureScorer      (src/Mm/GaussDiagonalMaximumFeatureScorer.hh:31-92)
 *    gpu-diagonal-sum            Mm::GaussDiagonalSumFeatureScorer          (src/Mm/GaussDiagonalMaximumFeatureScorer.hh:96-115)
 *    gpu-SIMD-diagonal-maximum   Mm::SimdGaussDiagonalMaximumFeatureScorer  (src/Mm/SimdFeatureScorer.hh:35-139)
 *    gpu-batch-diagonal-maximum-int  Mm::BatchIntFeatureScorer              (src/Mm/BatchFeatureScorer.hh:292-340)
 *    gpu-batch-diagonal-maximum-fast   Mm::BatchUnrolledIntFeatureScorer    (src/Mm/BatchFeatureScorer.hh:362-375)
 *    gpu-batch-diagonal-maximum-float  Mm::BatchFloatFeatureScorer          (src/Mm/BatchFeatureScorer.hh:226-290)
 *
 *  They implement the buffered protocol of Mm::FeatureScorer (src/Mm/FeatureScorer.hh:86-134) so that
 *  Speech::OfflineRecognizer / DelayedRecognizer / AlignmentNode / FeatureScorerNode and Flf's recognizer feed
 *  whole batches to the GPU with no caller change: frames are collected, the first score request for a frame
 *  without scores sends every pending frame through ONE kernel launch, and score(e) afterwards is an array read.
 *  buffer-size = 0 (default) is segment mode (never "filled", precedent src/Nn/TrainerFeatureScorer.cc:220-230);
 *  buffer-size = n is a ring of n frames with the state machine of src/Mm/BatchFeatureScorer.cc:40-105.
 */
#ifndef _MM_GPU_FEATURE_SCORER_HH
#define _MM_GPU_FEATURE_SCORER_HH

#include <deque>
#include <vector>

#include <Core/Parameter.hh>
#include <Core/ReferenceCounting.hh>
#include <Mm/AssigningFeatureScorer.hh>
#include <Mm/MixtureFeatureScorerElement.hh>
#include <Mm/MixtureSet.hh>

struct gmm_handle;

namespace MmGpu {

class GpuAligner;

class GpuFeatureScorer : public Mm::AssigningFeatureScorer {
    typedef Mm::AssigningFeatureScorer Precursor;
    friend class GpuAligner;  // rasr/GpuAligner.hh: takes the frames behind the context scorers to the device itself

public:
    enum Kind { kindMaximum, kindSum, kindQuantized, kindBatchInt };

    static const Core::ParameterInt   paramBufferSize;          // "buffer-size", as BatchFeatureScorer.cc:28-29
    static const Core::ParameterInt   paramDevice;              // "device"
    static const Core::ParameterInt   paramKernel;              // "kernel": 0 auto, 1 direct, 2 tensor
    static const Core::ParameterFloat paramMixtureWeightScale;  // as GaussDiagonalMaximumFeatureScorer.cc:37-39
    static const Core::ParameterFloat paramGaussianScale;       // as GaussDiagonalMaximumFeatureScorer.cc:41-43
    static const Core::ParameterBool  paramDensityAssignment;   // "with-density-assignment": training consumers
    static const Core::ParameterBool  paramDeferredScoring;     // "deferred-scoring": nothing is scored before it is asked for

protected:
    /** Page-locked host buffers, recycled: cudaMallocHost is a page-locking system call (milliseconds for the 18 MB of
     *  a 1000-frame segment), so blocks return their memory here instead of freeing it.  Shared by the scorer and its
     *  blocks, which may outlive it. */
    class PinnedPool : public Core::ReferenceCounted {
        struct Buffer {
            void*  ptr;
            size_t bytes;
        };
        std::vector<Buffer> free_;

    public:
        ~PinnedPool();
        void* acquire(size_t bytes, size_t& capacity);
        void  release(void* ptr, size_t capacity);
    };

    /** One GPU batch in page-locked host memory: the frames as they are handed in, and, once scored, their scores.
     *  Kept alive by the contexts that point into it (callers keep Scorer handles of a whole segment,
     *  src/Speech/AlignmentNode.cc:281-306).  Frames are written straight into the page-locked block the H2D copy
     *  reads, and contexts point at their block from the start: scoring a batch touches no context. */
    struct ScoreBlock : public Core::ReferenceCounted {
        Core::Ref<PinnedPool> pool;
        Mm::Score*            scores;    // [nFrames x ld]; 0 until the batch went to the GPU
        u32*                  best;      // [nFrames x ld]; 0 until a density assignment was requested
        float*                features;  // [capacity x dimension]
        size_t                scoresCap, bestCap, featuresCap;
        u32                   nFrames, capacity, dimension, ld;
        u64                   token;     // != 0: an asynchronous call is still filling the block (gmm_wait)
        ScoreBlock(Core::Ref<PinnedPool> pool, u32 capacity, u32 M, u32 D);
        ~ScoreBlock();
        bool append(const Mm::FeatureVector& f);
        void allocateScores();
        void allocateBest();
    };

    class Context : public AssigningContextScorer {
        friend class GpuFeatureScorer;
        friend class GpuAligner;
        const GpuFeatureScorer* parent_;
        Core::Ref<ScoreBlock>   block_;
        u32                     row_;
        void ready(bool withBest) const;

    public:
        Context(const GpuFeatureScorer* parent, Core::Ref<ScoreBlock> block, u32 row)
                : parent_(parent), block_(block), row_(row) {}
        virtual ~Context() {}
        virtual Mm::EmissionIndex    nEmissions() const;
        virtual Mm::Score            score(Mm::EmissionIndex e) const;
        virtual Mm::DensityInMixture bestDensity(Mm::EmissionIndex e) const;
        virtual Mm::Score            score(Mm::EmissionIndex e, Mm::DensityIndex dnsInMix) const;
        virtual void getDensityPosteriorProbabilities(Mm::EmissionIndex e, std::vector<Mm::Weight>& result) const;
    };
    friend class Context;

public:
    /** restrictions the reference's batch scorers put on the model (criticalError otherwise, as they do) */
    enum { pooledCovarianceOnly = 1, atMost48PaddedComponents = 2 };
    GpuFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> mixtureSet, Kind kind, u32 restrictions = 0);
    virtual ~GpuFeatureScorer();

    virtual Mm::EmissionIndex nMixtures() const {
        return nMixtures_;
    }
    virtual Mm::ComponentIndex dimension() const {
        return dimension_;
    }
    /** the training consumers' view of the model, as GaussDiagonalMaximumFeatureScorer.hh:79-87 answers it
     *  (wrapped e.g. by Mm::StatePosteriorFeatureScorer, StatePosteriorFeatureScorer.hh:205-236) */
    virtual Mm::DensityIndex nDensities() const {
        return nModelDensities_;
    }
    virtual const Mm::MixtureFeatureScorerElement& mixture(Mm::MixtureIndex m) const {
        return mixtureTable_[m];
    }
    virtual const std::vector<Mm::DensityIndex>& densitiesInMixture(Mm::MixtureIndex m) const {
        return mixtureTable_[m].densityIndices();
    }

    using Precursor::getAssigningScorer;
    /** add f, return the scorer of the oldest buffered frame (BatchFeatureScorer.cc:77-87) */
    virtual AssigningScorer getAssigningScorer(const Mm::FeatureVector& f) const;

    virtual void reset() const;
    virtual bool isBuffered() const {
        return true;
    }
    virtual void addFeature(const Mm::FeatureVector& f) const;
    virtual void addFeature(Core::Ref<const Mm::Feature> f) const {
        addFeature(*f->mainStream());
    }
    virtual Scorer flush() const;
    virtual bool   bufferFilled() const {
        return bufferSize_ > 0 && fifo_.size() >= size_t(bufferSize_ - 1);
    }
    virtual bool bufferEmpty() const {
        return fifo_.empty();
    }
    virtual u32 bufferSize() const {
        return bufferSize_ > 0 ? bufferSize_ : Core::Type<u32>::max;
    }
    /** segment mode keeps every frame of the current segment addressable by time (FeatureScorer.hh:157-163) */
    virtual bool hasTimeIndexedCache() const {
        return bufferSize_ == 0;
    }
    virtual Scorer getTimeIndexedScorer(u32 time) const;

    u32 nBatches() const {
        return nBatches_;
    }
    /** second passes over a batch for a density assignment that was not computed with its scores */
    u32 nAssignmentPasses() const {
        return nAssignmentPasses_;
    }
    /** the engine behind this scorer (include/gmm_b200.h), for callers that time or share it */
    gmm_handle* handle() const {
        return handle_;
    }

private:
    void initFloat(const Mm::MixtureSet& mixtureSet);
    void initQuantized(const Mm::MixtureSet& mixtureSet);
    void initBatchInt(const Mm::MixtureSet& mixtureSet);
    bool quantized() const {
        return kind_ == kindQuantized || kind_ == kindBatchInt;
    }
    void push(const Mm::FeatureVector& f) const;
    void scoreBlock(ScoreBlock& block, bool withBest, bool async = false) const;
    void assignDensities(ScoreBlock& block) const;
    void complete(ScoreBlock& block) const;
    void densityPosteriors(const ScoreBlock& block, u32 row, Mm::EmissionIndex e, std::vector<Mm::Weight>& result) const;
    void check(int rc, const char* what) const;

    enum { kSegmentBatch = 1024 };  // segment mode: frames per asynchronous batch
    Kind               kind_;
    gmm_handle*        handle_;
    u32                bufferSize_;
    bool               withBest_;
    bool               deferred_;
    Core::Ref<PinnedPool> pool_;
    Mm::EmissionIndex  nMixtures_;
    Mm::ComponentIndex dimension_;
    Mm::DensityIndex   nDensities_;       // (mixture, density) slots = columns of the kernels
    Mm::DensityIndex   nModelDensities_;  // distinct densities of the mixture set (shared ones once)
    std::vector<Mm::MixtureFeatureScorerElement> mixtureTable_;
    u32                maxDensities_;  // most densities in one mixture
    std::vector<u32>   offsets_;

    typedef Core::Ref<const Context>       ContextRef;
    mutable std::deque<ContextRef>         fifo_;     // frames handed in, oldest first, not yet handed out
    mutable Core::Ref<ScoreBlock>          staging_;  // frames handed in since the last batch (no scores yet)
    mutable std::vector<ContextRef>        segment_;  // segment mode: every frame since reset()
    mutable u32                            nBatches_, nAssignmentPasses_;
};

class GpuGaussDiagonalMaximumFeatureScorer : public GpuFeatureScorer {
public:
    GpuGaussDiagonalMaximumFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> m)
            : Core::Component(c), GpuFeatureScorer(c, m, kindMaximum) {}
};

class GpuGaussDiagonalSumFeatureScorer : public GpuFeatureScorer {
public:
    GpuGaussDiagonalSumFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> m)
            : Core::Component(c), GpuFeatureScorer(c, m, kindSum) {}
};

class GpuSimdGaussDiagonalMaximumFeatureScorer : public GpuFeatureScorer {
public:
    GpuSimdGaussDiagonalMaximumFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> m)
            : Core::Component(c), GpuFeatureScorer(c, m, kindQuantized) {}
};

/** The u8 kernel with the constants and the final scale of Mm::BatchIntFeatureScorer; pooled covariance only. */
class GpuBatchIntFeatureScorer : public GpuFeatureScorer {
public:
    GpuBatchIntFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> m)
            : Core::Component(c), GpuFeatureScorer(c, m, kindBatchInt) {}
};

/** Mm::BatchUnrolledIntFeatureScorer (`batch-diagonal-maximum-fast`, BatchFeatureScorer.cc:537-604): the numbers of the
 *  batch-int scorer (same init) from a loop unrolled for at most 48 padded components -- the same restriction here. */
class GpuBatchUnrolledIntFeatureScorer : public GpuFeatureScorer {
public:
    GpuBatchUnrolledIntFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> m)
            : Core::Component(c), GpuFeatureScorer(c, m, kindBatchInt, atMost48PaddedComponents) {}
};

/** Mm::BatchFloatFeatureScorer (`batch-diagonal-maximum-float`, BatchFeatureScorer.cc:138-234): 0.5 min_k(L - 2 logw +
 *  |mu isr - x isr|^2) over whitened operands, pooled covariance only -- the float maximum scorer (whose tensor kernel
 *  contracts the same whitened operands) within the float tolerance; weight / gaussian scales are not applied there
 *  and must be left at 1 here. */
class GpuBatchFloatFeatureScorer : public GpuFeatureScorer {
public:
    GpuBatchFloatFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> m)
            : Core::Component(c), GpuFeatureScorer(c, m, kindMaximum, pooledCovarianceOnly) {}
};

}  // namespace MmGpu

#endif  // _MM_GPU_FEATURE_SCORER_HH
