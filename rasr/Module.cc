#include "Module.hh"

#include <Mm/FeatureScorerFactory.hh>
#include <Mm/MixtureSetLoader.hh>
#include <Mm/Module.hh>

#include "GpuFeatureScorer.hh"

using namespace MmGpu;

Module_::Module_() {
    Mm::FeatureScorerFactory* factory = Mm::Module::instance().featureScorerFactory();
    // the names become legal values of `feature-scorer-type` (src/Mm/Module.cc:78-80)
    factory->registerFeatureScorer<GpuGaussDiagonalMaximumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(
            gpuDiagonalMaximum, "gpu-diagonal-maximum");
    factory->registerFeatureScorer<GpuGaussDiagonalSumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(
            gpuDiagonalSum, "gpu-diagonal-sum");
    factory->registerFeatureScorer<GpuSimdGaussDiagonalMaximumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(
            gpuSimdDiagonalMaximum, "gpu-SIMD-diagonal-maximum");
    factory->registerFeatureScorer<GpuBatchIntFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(
            gpuBatchIntDiagonalMaximum, "gpu-batch-diagonal-maximum-int");
    factory->registerFeatureScorer<GpuBatchUnrolledIntFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(
            gpuBatchIntDiagonalMaximumFast, "gpu-batch-diagonal-maximum-fast");
    factory->registerFeatureScorer<GpuBatchFloatFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(
            gpuBatchFloatDiagonalMaximum, "gpu-batch-diagonal-maximum-float");
}
