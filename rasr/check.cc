// TEST HARNESS: the GPU scorers and the reference's own scorers side by side INSIDE a RASR process, created through
// Mm::FeatureScorerFactory exactly as Mm::Module_::createFeatureScorer does (src/Mm/Module.cc:301-309) and driven by
// the buffered caller loop of Speech::AlignmentNode (src/Speech/AlignmentNode.cc:283-306).  Links the unmodified
// reference objects built by oracle/ref_build/Makefile; lives next to them in oracle/_ref/ (not product code).
//
//   rasr_gpu_check model=<gms> features=<f32> [buffer-size=N]
// exit code 0 iff: quantized scorer bit-identical, float direct kernel bit-identical, tensor kernel and sum scorer
// within 1e-4 relative, density assignments equal (float: up to ties within tolerance).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>

#include <Core/Application.hh>
#include <Mm/FeatureScorerFactory.hh>
#include <Mm/GaussDiagonalMaximumFeatureScorer.hh>
#include <Mm/MixtureSet.hh>
#include <Mm/MixtureSetLoader.hh>
#include <Mm/SimdFeatureScorer.hh>

#include "GpuFeatureScorer.hh"
#include "Module.hh"

// Harness stub (like oracle/ref_build/stubs): the factory only stores a loader object per scorer id; the file loaders
// themselves (src/Mm/MixtureSetLoader.cc -> Mm::Module_::readMixtureSet) pull in the whole estimator stack and are
// not needed here because the mixture set is built in memory.
Core::Ref<Mm::AbstractMixtureSet> Mm::AbstractMixtureSetLoader::load(const std::string&, const Core::Configuration&) const {
    return Core::Ref<Mm::AbstractMixtureSet>();
}

namespace {

template<class T>
bool readArray(FILE* f, std::vector<T>& v, size_t n) {
    v.resize(n);
    return n == 0 || fread(&v[0], sizeof(T), n, f) == n;
}

// "GMS1" container (rasr_b200/gms.py) -> Mm::MixtureSet
Core::Ref<Mm::MixtureSet> loadGms(const std::string& path) {
    Core::Ref<Mm::MixtureSet> ms;
    FILE*                     f = fopen(path.c_str(), "rb");
    if (!f)
        return ms;
    char             magic[4];
    u32              h[5];
    std::vector<u32> offs, mixDns, dMean, dCov;
    std::vector<f64> logw;
    std::vector<f32> means, vars;
    bool             ok = fread(magic, 1, 4, f) == 4 && !memcmp(magic, "GMS1", 4) && fread(h, 4, 5, f) == 5;
    const u32        D = h[0], M = h[1], nD = h[2], nM = h[3], nC = h[4];
    ok                 = ok && readArray(f, offs, M + 1);
    const size_t G     = ok ? offs[M] : 0;
    ok = ok && readArray(f, mixDns, G) && readArray(f, logw, G) && readArray(f, dMean, nD) && readArray(f, dCov, nD) &&
         readArray(f, means, size_t(nM) * D) && readArray(f, vars, size_t(nC) * D);
    fclose(f);
    if (!ok)
        return ms;
    ms = Core::Ref<Mm::MixtureSet>(new Mm::MixtureSet);
    ms->setDimension(D);
    for (u32 i = 0; i < nM; ++i) {
        Mm::Mean* mean = new Mm::Mean(D);
        std::copy(means.begin() + size_t(i) * D, means.begin() + size_t(i + 1) * D, mean->begin());
        ms->addMean(i, mean);
    }
    for (u32 c = 0; c < nC; ++c)
        ms->addCovariance(c, new Mm::DiagonalCovariance(std::vector<Mm::VarianceType>(vars.begin() + size_t(c) * D,
                                                                                      vars.begin() + size_t(c + 1) * D)));
    for (u32 d = 0; d < nD; ++d)
        ms->addDensity(d, new Mm::GaussDensity(dMean[d], dCov[d]));
    for (u32 m = 0; m < M; ++m) {
        Mm::Mixture* mix = new Mm::Mixture;
        for (u32 k = offs[m]; k < offs[m + 1]; ++k)
            mix->addLogDensity(mixDns[k], logw[k]);
        ms->addMixture(m, mix);
    }
    return ms;
}

struct Result {
    std::vector<f32> score;
    std::vector<u32> best;
};

// the caller loop of Speech/AlignmentNode.cc:283-306, then eager evaluation of every emission
Result drive(const Mm::FeatureScorer& fs, const std::vector<Mm::FeatureVector>& X, bool assigning) {
    std::vector<Mm::FeatureScorer::Scorer> scorers;
    fs.reset();
    for (size_t t = 0; t < X.size(); ++t) {
        if (fs.isBuffered() && !fs.bufferFilled())
            fs.addFeature(X[t]);
        else
            scorers.push_back(fs.getScorer(X[t]));
    }
    while (!fs.bufferEmpty())
        scorers.push_back(fs.flush());
    Result    r;
    const u32 M = fs.nMixtures();
    for (size_t t = 0; t < scorers.size(); ++t)
        for (u32 m = 0; m < M; ++m)
            r.score.push_back(scorers[t]->score(m));
    // density assignment through the unbuffered assigning interface (AssigningFeatureScorer.hh:66-70), as the
    // estimators use it (src/Mm/AbstractMixtureSetEstimator.cc:377-393)
    const Mm::AssigningFeatureScorer* afs = assigning ? dynamic_cast<const Mm::AssigningFeatureScorer*>(&fs) : 0;
    if (afs) {
        fs.reset();
        for (size_t t = 0; t < X.size(); ++t) {
            Mm::AssigningFeatureScorer::AssigningScorer a = afs->getAssigningScorer(X[t]);
            for (u32 m = 0; m < M; ++m)
                r.best.push_back(a->bestDensity(m));
        }
    }
    return r;
}

}  // namespace

class CheckApp : public Core::Application {
public:
    CheckApp() {
        setTitle("rasr-gpu-check");
        setDefaultLoadConfigurationFile(false);
        setDefaultOutputXmlHeader(false);
    }
    std::string getUsage() const {
        return "rasr_gpu_check model=<gms> features=<f32> [buffer-size=N]\n";
    }

    Mm::FeatureScorer* create(const Mm::FeatureScorerFactory& factory, u32 id, const std::string& sel, int kernel,
                              int bufferSize, Core::Ref<const Mm::MixtureSet> ms) {
        Core::Configuration c(config, sel);
        c.set(c.getSelection() + ".kernel", Core::form("%d", kernel));
        c.set(c.getSelection() + ".buffer-size", Core::form("%d", bufferSize));
        Mm::FeatureScorer* fs = factory.createFeatureScorer(id, c, ms);
        if (!fs)
            criticalError("Cannot create feature scorer!");  // as Mm/Module.cc:303-307
        return fs;
    }

    int main(const std::vector<std::string>& arguments) {
        std::map<std::string, std::string> args;
        for (size_t i = 0; i < arguments.size(); ++i) {
            size_t eq = arguments[i].find('=');
            if (eq != std::string::npos)
                args[arguments[i].substr(0, eq)] = arguments[i].substr(eq + 1);
        }
        Core::Ref<Mm::MixtureSet> msw = loadGms(args["model"]);
        if (!msw) {
            std::cerr << getUsage();
            return 2;
        }
        Core::Ref<const Mm::MixtureSet> ms(msw);
        std::ifstream                   in(args["features"].c_str(), std::ios::binary | std::ios::ate);
        if (!in) {
            std::cerr << getUsage();
            return 2;
        }
        std::vector<f32> flat(size_t(in.tellg()) / 4);
        in.seekg(0);
        in.read(reinterpret_cast<char*>(&flat[0]), flat.size() * 4);
        const u32                      D = ms->dimension();
        std::vector<Mm::FeatureVector> X(flat.size() / D, Mm::FeatureVector(D));
        for (size_t t = 0; t < X.size(); ++t)
            std::copy(flat.begin() + t * D, flat.begin() + (t + 1) * D, X[t].begin());
        const int bufferSize = args.count("buffer-size") ? atoi(args["buffer-size"].c_str()) : 0;

        // the factory, filled the way Mm::Module_::Module_ and MmGpu::Module_::Module_ fill it
        Mm::FeatureScorerFactory factory;
        enum { refMax = 1, refSum = 2, refSimd = 3 };
        factory.registerFeatureScorer<Mm::GaussDiagonalMaximumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(refMax, "diagonal-maximum");
        factory.registerFeatureScorer<Mm::GaussDiagonalSumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(refSum, "diagonal-sum");
        factory.registerFeatureScorer<Mm::SimdGaussDiagonalMaximumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(refSimd, "SIMD-diagonal-maximum");
        typedef MmGpu::Module_ G;
        factory.registerFeatureScorer<MmGpu::GpuGaussDiagonalMaximumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(G::gpuDiagonalMaximum, "gpu-diagonal-maximum");
        factory.registerFeatureScorer<MmGpu::GpuGaussDiagonalSumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(G::gpuDiagonalSum, "gpu-diagonal-sum");
        factory.registerFeatureScorer<MmGpu::GpuSimdGaussDiagonalMaximumFeatureScorer, Mm::MixtureSet, Mm::AbstractMixtureSetLoader>(G::gpuSimdDiagonalMaximum, "gpu-SIMD-diagonal-maximum");

        int failures = 0;
        struct Case {
            const char* name;
            u32         refId, gpuId;
            int         kernel;
            bool        exact;
        } cases[] = {{"diagonal-maximum / direct kernel", refMax, G::gpuDiagonalMaximum, 1, true},
                     {"diagonal-maximum / auto kernel", refMax, G::gpuDiagonalMaximum, 0, false},
                     {"diagonal-sum / auto kernel", refSum, G::gpuDiagonalSum, 0, false},
                     {"SIMD-diagonal-maximum", refSimd, G::gpuSimdDiagonalMaximum, 0, true}};
        for (size_t ci = 0; ci < sizeof(cases) / sizeof(cases[0]); ++ci) {
            const Case&                    k = cases[ci];
            Core::Ref<Mm::FeatureScorer>   ref(create(factory, k.refId, "ref", 0, 0, ms));
            Core::Ref<Mm::FeatureScorer>   gpu(create(factory, k.gpuId, "gpu", k.kernel, bufferSize, ms));
            Result                         a = drive(*ref, X, true), b = drive(*gpu, X, true);
            size_t                         nBits = 0, nBest = 0;
            double                         maxRel = 0.0;
            bool                           shape  = a.score.size() == b.score.size() && a.best.size() == b.best.size();
            for (size_t i = 0; shape && i < a.score.size(); ++i) {
                if (memcmp(&a.score[i], &b.score[i], 4))
                    ++nBits;
                if (std::isfinite(a.score[i]) && std::fabs(a.score[i]) < 1e37)
                    maxRel = std::max(maxRel, (double)std::fabs(a.score[i] - b.score[i]) / std::max(1.0, (double)std::fabs(a.score[i])));
                if (a.best[i] != b.best[i])
                    ++nBest;
            }
            bool ok = shape && (k.exact ? (nBits == 0 && nBest == 0) : (maxRel <= 1e-4 && nBest * 1000 <= a.best.size()));
            printf("%-36s %s  frames=%zu values=%zu differing-bits=%zu max-rel=%.3e best-mismatch=%zu batches=%u\n", k.name,
                   ok ? "PASS" : "FAIL", X.size(), a.score.size(), nBits, maxRel, nBest,
                   dynamic_cast<const MmGpu::GpuFeatureScorer*>(gpu.get())->nBatches());
            failures += !ok;
        }
        return failures ? 1 : 0;
    }
};

APPLICATION(CheckApp)
