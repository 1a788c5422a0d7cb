/** Module MM_GPU: registers the GPU GMM feature scorers in Mm::FeatureScorerFactory, the way src/Nn/Module.cc:47-65
 *  registers the NN scorers.  Enabled by `MODULES += MODULE_MM_GPU` in Modules.make; tools add
 *  `INIT_MODULE(MmGpu)` next to INIT_MODULE(Mm) (src/Tools/SpeechRecognizer/SpeechRecognizer.cc:46-56).
 */
#ifndef _MM_GPU_MODULE_HH
#define _MM_GPU_MODULE_HH

#include <Core/Singleton.hh>

namespace MmGpu {

class Module_ {
public:
    Module_();
    /** ids in a private range, like Nn's 0x300 (src/Nn/Module.hh:36-45); Mm's own are src/Mm/Module.hh:48-70 */
    enum { FeatureScorerTypeOffset = 0x400 };
    enum FeatureScorerType {
        gpuDiagonalMaximum     = FeatureScorerTypeOffset,
        gpuDiagonalSum         = FeatureScorerTypeOffset + 1,
        gpuSimdDiagonalMaximum = FeatureScorerTypeOffset + 2,
        gpuBatchIntDiagonalMaximum = FeatureScorerTypeOffset + 3,
        gpuBatchIntDiagonalMaximumFast = FeatureScorerTypeOffset + 4,
        gpuBatchFloatDiagonalMaximum   = FeatureScorerTypeOffset + 5,
    };
};

typedef Core::SingletonHolder<Module_> Module;

}  // namespace MmGpu

#endif  // _MM_GPU_MODULE_HH
