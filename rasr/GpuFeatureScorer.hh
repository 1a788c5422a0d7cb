/** RASR-side plugin: GPU GMM feature scorers behind Mm::FeatureScorerFactory.
 *
 *  Compiled INTO RASR (module MODULE_MM_GPU, see rasr/Module.hh and INTEGRATION.md); binds only to the C-ABI of
 *  include/gmm_b200.h (libgmm_b200.so, hand-written CUDA for sm_100a).  Three scorers, drop-in for
 *
 *    gpu-diagonal-maximum        Mm::GaussDiagonalMaximumFeatureScorer      (src/Mm/GaussDiagonalMaximumFeatureScorer.hh:31-92)
 *    gpu-diagonal-sum            Mm::GaussDiagonalSumFeatureScorer          (src/Mm/GaussDiagonalMaximumFeatureScorer.hh:96-115)
 *    gpu-SIMD-diagonal-maximum   Mm::SimdGaussDiagonalMaximumFeatureScorer  (src/Mm/SimdFeatureScorer.hh:35-139)
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
#include <Mm/MixtureSet.hh>

struct gmm_handle;

namespace MmGpu {

class GpuFeatureScorer : public Mm::AssigningFeatureScorer {
    typedef Mm::AssigningFeatureScorer Precursor;

public:
    enum Kind { kindMaximum, kindSum, kindQuantized };

    static const Core::ParameterInt   paramBufferSize;          // "buffer-size", as BatchFeatureScorer.cc:28-29
    static const Core::ParameterInt   paramDevice;              // "device"
    static const Core::ParameterInt   paramKernel;              // "kernel": 0 auto, 1 direct, 2 tensor
    static const Core::ParameterFloat paramMixtureWeightScale;  // as GaussDiagonalMaximumFeatureScorer.cc:37-39
    static const Core::ParameterFloat paramGaussianScale;       // as GaussDiagonalMaximumFeatureScorer.cc:41-43

protected:
    /** Scores of one GPU batch in page-locked host memory; kept alive by the contexts that point into it
     *  (callers keep Scorer handles of a whole segment, src/Speech/AlignmentNode.cc:281-306). */
    struct ScoreBlock : public Core::ReferenceCounted {
        Mm::Score*         scores;
        u32*               best;  // 0 until a density assignment was requested
        std::vector<float> features;
        u32                nFrames, nMixtures;
        ScoreBlock(u32 T, u32 M);
        ~ScoreBlock();
    };

    class Context : public AssigningContextScorer {
        friend class GpuFeatureScorer;
        const GpuFeatureScorer*   parent_;
        mutable Core::Ref<ScoreBlock> block_;
        mutable u32               row_;
        mutable std::vector<float> pending_;  // the frame until its batch has been scored

    public:
        Context(const GpuFeatureScorer* parent, const Mm::FeatureVector& f);
        virtual ~Context() {}
        virtual Mm::EmissionIndex    nEmissions() const;
        virtual Mm::Score            score(Mm::EmissionIndex e) const;
        virtual Mm::DensityInMixture bestDensity(Mm::EmissionIndex e) const;
        virtual Mm::Score            score(Mm::EmissionIndex e, Mm::DensityIndex dnsInMix) const;
        virtual void getDensityPosteriorProbabilities(Mm::EmissionIndex e, std::vector<Mm::Weight>& result) const;
    };
    friend class Context;

public:
    GpuFeatureScorer(const Core::Configuration& c, Core::Ref<const Mm::MixtureSet> mixtureSet, Kind kind);
    virtual ~GpuFeatureScorer();

    virtual Mm::EmissionIndex nMixtures() const {
        return nMixtures_;
    }
    virtual Mm::ComponentIndex dimension() const {
        return dimension_;
    }
    virtual Mm::DensityIndex nDensities() const {
        return nDensities_;
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

private:
    void initFloat(const Mm::MixtureSet& mixtureSet);
    void initQuantized(const Mm::MixtureSet& mixtureSet);
    void push(const Mm::FeatureVector& f) const;
    void scorePending(bool withBest) const;
    void assignDensities(ScoreBlock& block) const;
    void check(int rc, const char* what) const;

    Kind               kind_;
    gmm_handle*        handle_;
    u32                bufferSize_;
    Mm::EmissionIndex  nMixtures_;
    Mm::ComponentIndex dimension_;
    Mm::DensityIndex   nDensities_;
    std::vector<u32>   offsets_;

    typedef Core::Ref<const Context>       ContextRef;
    mutable std::deque<ContextRef>         fifo_;     // frames handed in, oldest first, not yet handed out
    mutable std::vector<ContextRef>        pending_;  // frames without scores yet (time order)
    mutable std::vector<ContextRef>        segment_;  // segment mode: every frame since reset()
    mutable u32                            nBatches_;
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

}  // namespace MmGpu

#endif  // _MM_GPU_FEATURE_SCORER_HH
