// TEST HARNESS: the GPU scorers and the reference's own scorers side by side INSIDE a RASR process.  Both come out of
// the reference's own Mm::Module_ -- the factory MmGpu::Module_ registers in (rasr/Module.cc, as src/Nn/Module.cc:47-65
// does), selected by `feature-scorer-type` (src/Mm/Module.cc:78-80,301-309) and wrapped in Mm::FeatureScorerScaling
// (createScaledFeatureScorer, src/Mm/Module.cc:326-328), i.e. exactly what Am::ClassicAcousticModel::loadMixtureSet
// hands to Speech and Search (src/Am/ClassicAcousticModel.cc:153-165).  Links the unmodified reference objects built by
// oracle/ref_build/Makefile; lives next to them in oracle/_ref/ (not product code).
//
//   rasr_gpu_check model=<gms> features=<f32> [buffer-size=N]
// One line per case, PASS / FAIL; exit code 0 iff all pass:
//   quantized scorer and float direct kernel bit-identical to the reference, tensor kernel and sum scorer within 1e-4
//   relative, density assignments equal (float: up to ties within tolerance) -- all through the scaled wrapper and the
//   buffered caller loop of Speech::AlignmentNode (src/Speech/AlignmentNode.cc:283-306); then the delayed-recognizer
//   loop (src/Speech/DelayedRecognizer.cc:72-175), the time-indexed cache (src/Mm/FeatureScorer.hh:157-163), the sum
//   scorer's density posteriors (src/Mm/AssigningFeatureScorer.hh:46-47) and with-density-assignment; and a TIMING line.
#include <chrono>
#include <cstdarg>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <deque>
#include <fstream>
#include <iostream>
#include <map>

#include <Core/Application.hh>
#include <Mm/FeatureScorerFactory.hh>
#include <Mm/GaussDiagonalMaximumFeatureScorer.hh>
#include <Mm/MixtureSet.hh>
#include <Mm/MixtureSetLoader.hh>
#include <Mm/Module.hh>
#include <Mm/ScaledFeatureScorer.hh>
#include <Mm/SimdFeatureScorer.hh>

#include <Fsa/Static.hh>
#include <Search/Aligner.hh>
#include <Speech/Alignment.hh>

#include <gmm_b200.h>

#include <ref_table_am.hh>  // oracle/ref_build: an acoustic model whose state tying is a table (test double)

#include "GpuAligner.hh"
#include "GpuFeatureScorer.hh"
#include "Module.hh"

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

// the assigning context scorer's class name is protected; its public handle type names it
typedef std::remove_reference<decltype(*std::declval<Mm::AssigningFeatureScorer::AssigningScorer>())>::type AssigningContext;

struct Result {
    std::vector<f32> score;
    std::vector<u32> best;
};

// the caller loop of Speech/AlignmentNode.cc:283-306, then eager evaluation of every emission; `assign` (the scorer
// inside the scaling wrapper) answers the density assignment through the unbuffered assigning interface
// (AssigningFeatureScorer.hh:66-70), as the estimators use it (src/Mm/AbstractMixtureSetEstimator.cc:377-393)
Result drive(const Mm::FeatureScorer& fs, const Mm::FeatureScorer* assign, const std::vector<Mm::FeatureVector>& X) {
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
    const Mm::AssigningFeatureScorer* afs = dynamic_cast<const Mm::AssigningFeatureScorer*>(assign);
    if (afs) {
        afs->reset();
        for (size_t t = 0; t < X.size(); ++t) {
            Mm::AssigningFeatureScorer::AssigningScorer a = afs->getAssigningScorer(X[t]);
            for (u32 m = 0; m < M; ++m)
                r.best.push_back(a->bestDensity(m));
        }
    }
    return r;
}

// Speech::RecognizerDelayHandler (src/Speech/DelayedRecognizer.cc:72-175) without a cache and with look-ahead length 0:
// a feature buffer of scorer->bufferSize() frames, initializeBatchScorer() primes the scorer with addFeature() until
// bufferFilled(), every further frame goes in through getScorer(), the rest comes out through flush().
Result driveDelayed(const Mm::FeatureScorer& fs, const std::vector<Mm::FeatureVector>& X) {
    Result                        r;
    const u32                     M = fs.nMixtures();
    const size_t                  bufferSize = std::min<size_t>(fs.bufferSize(), X.size() + 1);
    std::deque<Mm::FeatureVector> buffer;
    size_t                        offset = 0;
    bool                          initialize = true;
    fs.reset();
    auto prime = [&]() {  // initializeBatchScorer
        size_t pos = 0;
        while (!fs.bufferFilled() && pos < buffer.size())
            fs.addFeature(buffer[pos++]);
        offset     = pos;
        initialize = false;
    };
    auto feed = [&](Mm::FeatureScorer::Scorer sc) {
        for (u32 m = 0; m < M; ++m)
            r.score.push_back(sc->score(m));
    };
    for (size_t t = 0; t < X.size(); ++t) {  // add()
        if (buffer.size() < bufferSize) {
            buffer.push_back(X[t]);
            continue;
        }
        if (fs.isBuffered() && initialize)
            prime();
        buffer.push_back(X[t]);
        Mm::FeatureScorer::Scorer sc = fs.getScorer(buffer[offset]);
        buffer.pop_front();
        feed(sc);
    }
    while (!buffer.empty()) {  // flush()
        if (fs.isBuffered() && initialize)
            prime();
        Mm::FeatureScorer::Scorer sc = buffer.size() > offset ? fs.getScorer(buffer[offset]) : fs.flush();
        buffer.pop_front();
        feed(sc);
    }
    return r;
}

struct Diff {
    size_t nBits, nBest;
    double maxRel;
    bool   shape;
};

Diff compare(const Result& a, const Result& b) {
    Diff d = {0, 0, 0.0, a.score.size() == b.score.size() && a.best.size() == b.best.size()};
    for (size_t i = 0; d.shape && i < a.score.size(); ++i) {
        if (memcmp(&a.score[i], &b.score[i], 4))
            ++d.nBits;
        if (std::isfinite(a.score[i]) && std::fabs(a.score[i]) < 1e37)
            d.maxRel = std::max(d.maxRel, (double)std::fabs(a.score[i] - b.score[i]) / std::max(1e-3, (double)std::fabs(a.score[i])));
    }
    for (size_t i = 0; d.shape && i < a.best.size(); ++i)
        if (a.best[i] != b.best[i])
            ++d.nBest;
    return d;
}

// the caller loop of Speech/AlignmentNode.cc:283-306 for the frames [t0, t1): the scorers a segment's aligner is fed
std::vector<Mm::FeatureScorer::Scorer> segmentScorers(const Mm::FeatureScorer& fs, const std::vector<Mm::FeatureVector>& X,
                                                      size_t t0, size_t t1) {
    std::vector<Mm::FeatureScorer::Scorer> scorers;
    fs.reset();
    for (size_t t = t0; t < t1; ++t) {
        if (fs.isBuffered() && !fs.bufferFilled())
            fs.addFeature(X[t]);
        else
            scorers.push_back(fs.getScorer(X[t]));
    }
    if (fs.isBuffered())
        while (!fs.bufferEmpty())
            scorers.push_back(fs.flush());
    return scorers;
}

// Left-to-right allophone-state acceptor with loop / forward / skip transitions over a label sequence: the shape
// Speech::AllophoneStateGraphBuilder gives a pronunciation; arcs carry LABELS (allophone states).
Fsa::ConstAutomatonRef linearModel(const std::vector<u32>& labels, f32 loop, f32 forward, f32 skip) {
    Fsa::StaticAutomaton* model = new Fsa::StaticAutomaton();
    model->setType(Fsa::TypeAcceptor);
    model->setSemiring(Fsa::TropicalSemiring);
    const u32 n = labels.size();
    for (u32 s = 0; s <= n; ++s) {
        Fsa::State* sp = new Fsa::State(s);
        if (s > 0)
            sp->newArc(s, Fsa::Weight(loop), labels[s - 1]);
        if (s < n)
            sp->newArc(s + 1, Fsa::Weight(forward), labels[s]);
        if (s + 1 < n)
            sp->newArc(s + 2, Fsa::Weight(skip), labels[s + 1]);
        if (s == n)
            sp->setFinal(Fsa::Weight(0.0));
        model->setState(sp);
    }
    model->setInitialStateId(0);
    return Fsa::ConstAutomatonRef(model);
}

struct AlignDiff {
    size_t items, differing;  // items of the reference; (time, label) pairs whose weight differs by more than tol
    double maxDiff;
    bool   scoreBits;
    double scoreRel;
};

// item sets per (time, label): the reference's order within a frame is that of an unstable sort on the weight
AlignDiff compareAlignments(const Speech::Alignment& a, const Speech::Alignment& b, double tol) {
    std::map<std::pair<u32, u32>, std::pair<double, double> > w;
    for (Speech::Alignment::const_iterator i = a.begin(); i != a.end(); ++i)
        w[std::make_pair(u32(i->time), u32(i->emission))].first += i->weight;
    for (Speech::Alignment::const_iterator i = b.begin(); i != b.end(); ++i)
        w[std::make_pair(u32(i->time), u32(i->emission))].second += i->weight;
    AlignDiff d = {a.size(), 0, 0.0, false, 0.0};
    for (std::map<std::pair<u32, u32>, std::pair<double, double> >::const_iterator i = w.begin(); i != w.end(); ++i) {
        const double diff = std::fabs(i->second.first - i->second.second);
        d.maxDiff         = std::max(d.maxDiff, diff);
        d.differing += diff > tol;
    }
    const f32 sa = a.score(), sb = b.score();
    d.scoreBits = !memcmp(&sa, &sb, 4);
    d.scoreRel  = std::fabs(double(sa) - double(sb)) / std::max(1.0, std::fabs(double(sa)));
    return d;
}

double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

class CheckApp : public Core::Application {
    int failures_;

    void report(const char* name, bool ok, const char* fmt, ...) {
        char    buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        printf("%-44s %s  %s\n", name, ok ? "PASS" : "FAIL", buf);
        failures_ += !ok;
    }

public:
    CheckApp()
            : failures_(0) {
        setTitle("rasr-gpu-check");
        setDefaultLoadConfigurationFile(false);
        setDefaultOutputXmlHeader(false);
    }
    std::string getUsage() const {
        return "rasr_gpu_check model=<gms> features=<f32> [buffer-size=N]\n";
    }

    // `*.mixture-set.feature-scorer-type = <name>` and the scorer's own parameters, as a RASR config file would set them
    Core::Configuration scorerConfig(const std::string& sel, const std::string& type, int kernel, int bufferSize,
                                     bool withAssignment = false) {
        Core::Configuration c(config, sel);
        const std::string   s = c.getSelection();
        c.set(s + ".feature-scorer-type", type);
        c.set(s + ".kernel", Core::form("%d", kernel));
        c.set(s + ".buffer-size", Core::form("%d", bufferSize));
        c.set(s + ".with-density-assignment", withAssignment ? "true" : "false");
        c.set(s + ".scale", "0.75");  // Mc::Component::paramScale of the scaling wrapper
        return c;
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
        const u32                      D = ms->dimension(), M = ms->nMixtures();
        std::vector<Mm::FeatureVector> X(flat.size() / D, Mm::FeatureVector(D));
        for (size_t t = 0; t < X.size(); ++t)
            std::copy(flat.begin() + t * D, flat.begin() + (t + 1) * D, X[t].begin());
        const int bufferSize = args.count("buffer-size") ? atoi(args["buffer-size"].c_str()) : 0;

        // INIT_MODULE(Mm); INIT_MODULE(MmGpu);  (src/Tools/SpeechRecognizer/SpeechRecognizer.cc:46-56)
        Mm::Module::instance();
        MmGpu::Module::instance();
        const Mm::Module_& mm = Mm::Module::instance();

        // ---- 1-7: every scorer pair through factory + scaling wrapper + the AlignmentNode loop ----------------------
        struct Case {
            const char *name, *refType, *gpuType;
            int         kernel;
            bool        exact, pooledOnly;
        } cases[] = {{"diagonal-maximum / direct kernel", "diagonal-maximum", "gpu-diagonal-maximum", 1, true, false},
                     {"diagonal-maximum / auto kernel", "diagonal-maximum", "gpu-diagonal-maximum", 0, false, false},
                     {"diagonal-sum / auto kernel", 0, "gpu-diagonal-sum", 0, false, false},
                     {"SIMD-diagonal-maximum", "SIMD-diagonal-maximum", "gpu-SIMD-diagonal-maximum", 0, true, false},
                     // the reference's scorer is a plain buffered FeatureScorer without a density assignment
                     {"batch-diagonal-maximum-int", "batch-diagonal-maximum-int", "gpu-batch-diagonal-maximum-int", 0, true, true},
                     {"batch-diagonal-maximum-fast", "batch-diagonal-maximum-fast", "gpu-batch-diagonal-maximum-fast", 0, true, true},
                     {"batch-diagonal-maximum-float", "batch-diagonal-maximum-float", "gpu-batch-diagonal-maximum-float", 0, false, true}};
        for (size_t ci = 0; ci < sizeof(cases) / sizeof(cases[0]); ++ci) {
            const Case&                  k = cases[ci];
            if (k.pooledOnly && (ms->nCovariances() != 1 || (ci == 5 && D > 48))) {
                printf("%-44s SKIP  needs a globally pooled variance (criticalError in both, BatchFeatureScorer.cc:340-342)\n", k.name);
                continue;
            }
            // (the reference's batch scorers need a ring of at least one frame: their default buffer-size, 4)
            Core::Configuration          rc = scorerConfig("ref", k.refType ? k.refType : "diagonal-maximum", 0, k.pooledOnly ? 4 : 0);
            Core::Configuration          gc = scorerConfig("gpu", k.gpuType, k.kernel, bufferSize);
            // the reference does not register its sum scorer: constructed directly
            Core::Ref<Mm::FeatureScorer> ref = k.refType ? mm.createFeatureScorer(rc, ms)
                                                         : Core::Ref<Mm::FeatureScorer>(new Mm::GaussDiagonalSumFeatureScorer(rc, ms));
            Core::Ref<Mm::FeatureScorer> gpu = mm.createFeatureScorer(gc, ms);
            Core::Ref<Mm::ScaledFeatureScorer> refScaled = mm.createScaledFeatureScorer(rc, ref);
            Core::Ref<Mm::ScaledFeatureScorer> gpuScaled = mm.createScaledFeatureScorer(gc, gpu);
            // the reference's batch scorers keep the frames in a ring of buffer-size entries: a context scorer must
            // be read before the ring comes round, as the recognizers do -- the delayed loop reads at once
            Result a = k.pooledOnly ? driveDelayed(*refScaled, X) : drive(*refScaled, ref.get(), X);
            Result b = k.pooledOnly ? driveDelayed(*gpuScaled, X) : drive(*gpuScaled, gpu.get(), X);
            Diff   d = compare(a, b);
            bool   ok = d.shape && (k.exact ? (d.nBits == 0 && d.nBest == 0) : (d.maxRel <= 1e-4 && d.nBest * 1000 <= a.best.size()));
            // the wrapper multiplies by its scale (ScaledFeatureScorer.hh:62-64): spot-check against the bare scorer
            gpu->reset();
            Mm::FeatureScorer::Scorer bare = gpu->isBuffered() ? (gpu->addFeature(X[0]), gpu->flush()) : gpu->getScorer(X[0]);
            ok = ok && b.score[0] == Mm::Score(0.75) * bare->score(0);
            report(k.name, ok, "frames=%zu values=%zu differing-bits=%zu max-rel=%.3e best-mismatch=%zu batches=%u", X.size(),
                   a.score.size(), d.nBits, d.maxRel, d.nBest, dynamic_cast<const MmGpu::GpuFeatureScorer*>(gpu.get())->nBatches());
        }

        // ---- 5: the delayed-recognizer loop (look-ahead / Flf recognizers), streaming and segment mode ---------------
        {
            Core::Configuration          rc = scorerConfig("ref", "diagonal-maximum", 0, 0);
            Core::Ref<Mm::FeatureScorer> ref = mm.createFeatureScorer(rc, ms);
            Result                       a   = driveDelayed(*mm.createScaledFeatureScorer(rc, ref), X);
            bool                         ok  = true;
            double                       worst = 0.0;
            u32                          batches = 0;
            const int                    sizes[] = {bufferSize > 0 ? bufferSize : 16, 0};
            for (int i = 0; i < 2; ++i) {
                Core::Configuration          gc  = scorerConfig("gpu", "gpu-diagonal-maximum", 1, sizes[i]);
                Core::Ref<Mm::FeatureScorer> gpu = mm.createFeatureScorer(gc, ms);
                Diff                         d   = compare(a, driveDelayed(*mm.createScaledFeatureScorer(gc, gpu), X));
                ok      = ok && d.shape && d.nBits == 0;
                worst   = std::max(worst, d.maxRel);
                batches = dynamic_cast<const MmGpu::GpuFeatureScorer*>(gpu.get())->nBatches();
            }
            report("delayed-recognizer loop (direct kernel)", ok, "buffer sizes %d and 0: bit-identical=%d max-rel=%.3e batches(segment mode)=%u",
                   sizes[0], int(ok), worst, batches);
        }

        // ---- 6: time-indexed cache of the segment mode (SegmentwiseAlignmentGenerator.cc:90-156) ----------------------
        {
            Core::Configuration          gc  = scorerConfig("gpu", "gpu-diagonal-maximum", 0, 0);
            Core::Ref<Mm::FeatureScorer> gpu = mm.createFeatureScorer(gc, ms);
            bool                         ok  = gpu->hasTimeIndexedCache();
            gpu->reset();
            for (size_t t = 0; t < X.size(); ++t)
                gpu->addFeature(X[t]);
            std::vector<Mm::FeatureScorer::Scorer> byFlush;
            size_t                                 bad = 0;
            for (size_t t = X.size(); ok && t-- > 0;) {  // random access by time, backwards
                Mm::FeatureScorer::Scorer s = gpu->getTimeIndexedScorer(t);
                byFlush.push_back(s);
            }
            std::vector<Mm::FeatureScorer::Scorer> inOrder;
            while (ok && !gpu->bufferEmpty())
                inOrder.push_back(gpu->flush());
            for (size_t t = 0; ok && t < X.size(); ++t)
                for (u32 m = 0; m < M; m += 7)
                    bad += byFlush[X.size() - 1 - t]->score(m) != inOrder[t]->score(m);
            ok = ok && bad == 0 && inOrder.size() == X.size();
            report("time-indexed scorers (segment mode)", ok, "frames=%zu mismatches=%zu batches=%u", X.size(), bad,
                   dynamic_cast<const MmGpu::GpuFeatureScorer*>(gpu.get())->nBatches());
        }

        // ---- 7: density posteriors of the sum scorer (AssigningFeatureScorer.hh:46-47; consumer
        //         AbstractMixtureSetEstimator.cc:386-395) ---------------------------------------------------------------
        {
            Core::Configuration                     rc = scorerConfig("ref", "diagonal-maximum", 0, 0);
            Core::Configuration                     gc = scorerConfig("gpu", "gpu-diagonal-sum", 0, 0);
            Core::Ref<Mm::AssigningFeatureScorer>   ref(new Mm::GaussDiagonalSumFeatureScorer(rc, ms));
            Core::Ref<Mm::AssigningFeatureScorer>   gpu = mm.createAssigningFeatureScorer(gc, ms);
            double                                  worst = 0.0;
            size_t                                  n = 0, sizeBad = 0;
            // the reference's sum scorer caches density scores by context ADDRESS (GaussDiagonalMaximumFeatureScorer.cc:
            // 239-246): keep every context alive so that no address is reused with stale scores
            std::vector<Mm::AssigningFeatureScorer::AssigningScorer> keep;
            for (size_t t = 0; t < X.size(); t += 3) {
                Mm::AssigningFeatureScorer::AssigningScorer a = ref->getAssigningScorer(X[t]);
                Mm::AssigningFeatureScorer::AssigningScorer b = gpu->getAssigningScorer(X[t]);
                keep.push_back(a);
                for (int j = 0; j < 3; ++j) {
                    const Mm::EmissionIndex e = (t * 7 + j * 13) % M;
                    std::vector<Mm::Weight> pa, pb;
                    a->getDensityPosteriorProbabilities(e, pa);
                    b->getDensityPosteriorProbabilities(e, pb);
                    sizeBad += pa.size() != pb.size();
                    for (size_t k = 0; k < pa.size() && k < pb.size(); ++k, ++n)
                        worst = std::max(worst, std::fabs(pa[k] - pb[k]) / std::max(1e-6, std::fabs(pa[k])));
                }
            }
            report("density posteriors (sum scorer)", sizeBad == 0 && n > 0 && worst <= 1e-4, "values=%zu max-rel=%.3e size-mismatch=%zu",
                   n, worst, sizeBad);
        }

        // ---- the model accessors of the assigning interface (GaussDiagonalMaximumFeatureScorer.hh:79-87) -----------------
        {
            Core::Configuration                   rc  = scorerConfig("ref", "diagonal-maximum", 0, 0);
            Core::Configuration                   gc  = scorerConfig("gpu", "gpu-diagonal-maximum", 0, 0);
            Core::Ref<Mm::AssigningFeatureScorer> ref = mm.createAssigningFeatureScorer(rc, ms);
            Core::Ref<Mm::AssigningFeatureScorer> gpu = mm.createAssigningFeatureScorer(gc, ms);
            size_t bad = (ref->nDensities() != gpu->nDensities()) + (ref->dimension() != gpu->dimension()) +
                         (ref->nMixtures() != gpu->nMixtures());
            for (u32 m = 0; m < M; ++m) {
                bad += ref->densitiesInMixture(m) != gpu->densitiesInMixture(m);
                bad += ref->mixture(m).minus2LogWeights() != gpu->mixture(m).minus2LogWeights();
            }
            report("nDensities / mixture(m) / densitiesInMixture(m)", bad == 0, "mismatches=%zu densities=%u", bad,
                   u32(gpu->nDensities()));
        }

        // ---- 8: with-density-assignment: score AND bestDensity of buffered frames from ONE pass -----------------------
        {
            u32    passes[2], batches[2];
            Result r[2];
            for (int i = 0; i < 2; ++i) {
                Core::Configuration          gc  = scorerConfig("gpu", "gpu-diagonal-maximum", 1, 0, i == 1);
                Core::Ref<Mm::FeatureScorer> gpu = mm.createFeatureScorer(gc, ms);
                gpu->reset();
                for (size_t t = 0; t < X.size(); ++t)
                    gpu->addFeature(X[t]);
                while (!gpu->bufferEmpty()) {
                    Mm::FeatureScorer::Scorer s  = gpu->flush();
                    const AssigningContext*   as = dynamic_cast<const AssigningContext*>(s.get());
                    for (u32 m = 0; as && m < M; m += 5) {
                        r[i].score.push_back(as->score(m));  // score first, then the assignment, as
                        r[i].best.push_back(as->bestDensity(m));  // AbstractMixtureSetEstimator.cc:377-393 asks
                    }
                }
                const MmGpu::GpuFeatureScorer* g = dynamic_cast<const MmGpu::GpuFeatureScorer*>(gpu.get());
                passes[i]  = g->nAssignmentPasses();
                batches[i] = g->nBatches();
            }
            Diff d = compare(r[0], r[1]);
            report("with-density-assignment", d.shape && !r[0].best.empty() && d.nBits == 0 && d.nBest == 0 && batches[1] == 1 && passes[1] == 0,
                   "identical=%d; GPU passes: %u with, %u without", int(d.nBits == 0 && d.nBest == 0), batches[1] + passes[1],
                   batches[0] + passes[0]);
        }

        // ---- 9: MmGpu::GpuAligner against the reference's own Search::Aligner, inside this process --------------------
        // Segments of the feature file, each with a left-to-right acceptor over random allophone-state LABELS; the
        // acoustic model ties 3 labels to every mixture, so parallel labels of one mixture occur (they must stay apart in
        // the items).  The reference aligner runs unpruned on the scorers of the SAME engine through the scaling wrapper
        // (scale 0.75): it reads the full score matrix from the host; the GPU aligner gets the same scorers from a
        // deferred-scoring instance and must never cause a full scoring pass.  Viterbi: items and score bit for bit
        // (kernel 1 = direct: also against the reference's own CPU scorer).  Baum-Welch: weights within 1e-4.
        {
            const u32        nLabels = 3 * M;
            std::vector<u32> tie(nLabels);
            for (u32 l = 0; l < nLabels; ++l)
                tie[l] = l % M;  // labels l, l + M, l + 2M share mixture l
            Core::Ref<const Am::AcousticModel> am(new TableAcousticModel(select("acoustic-model"), M, tie));
            Core::Configuration                ac(config, "aligner");
            ac.set(ac.getSelection() + ".min-acoustic-pruning", "1e30");  // one unpruned pass (Aligner.cc:584-611)
            ac.set(ac.getSelection() + ".max-acoustic-pruning", "1e30");
            const size_t segLen[] = {1, 0, 37, 120, 250, 400};
            for (int mode = 0; mode < 2; ++mode)
                for (int kernel = 1; kernel >= 0; --kernel) {
                    Core::Configuration gc = scorerConfig("gpu", "gpu-diagonal-maximum", kernel, 0);
                    Core::Configuration dc = scorerConfig("gpu-deferred", "gpu-diagonal-maximum", kernel, 0);
                    dc.set(dc.getSelection() + ".deferred-scoring", "true");
                    Core::Configuration rc = scorerConfig("ref", "diagonal-maximum", 0, 0);
                    Core::Ref<Mm::FeatureScorer>       gpu = mm.createFeatureScorer(gc, ms), dfr = mm.createFeatureScorer(dc, ms);
                    Core::Ref<Mm::ScaledFeatureScorer> gpuScaled = mm.createScaledFeatureScorer(gc, gpu);
                    Core::Ref<Mm::ScaledFeatureScorer> dfrScaled = mm.createScaledFeatureScorer(dc, dfr);
                    Core::Ref<Mm::ScaledFeatureScorer> cpuScaled = mm.createScaledFeatureScorer(rc, mm.createFeatureScorer(rc, ms));
                    Search::Aligner   refAligner(ac), cpuAligner(ac);
                    MmGpu::GpuAligner gpuAligner(ac);
                    refAligner.selectMode(mode ? Search::Aligner::modeBaumWelch : Search::Aligner::modeViterbi);
                    cpuAligner.selectMode(mode ? Search::Aligner::modeBaumWelch : Search::Aligner::modeViterbi);
                    gpuAligner.selectMode(mode ? MmGpu::GpuAligner::modeBaumWelch : MmGpu::GpuAligner::modeViterbi);
                    size_t  t0 = 0, nItems = 0, nDiffer = 0, nCpuDiffer = 0, nScoreBits = 0, nSeg = 0, nReached = 0;
                    double  maxDiff = 0, maxRel = 0, secRef = 0, secGpu = 0;
                    u32     seed = 12345u + 17u * kernel + mode;
                    for (size_t si = 0; si < sizeof(segLen) / sizeof(segLen[0]) && t0 + segLen[si] <= X.size(); ++si) {
                        const size_t     T = segLen[si];
                        std::vector<u32> labels(std::max<size_t>(1, T / 3));
                        for (size_t i = 0; i < labels.size(); ++i) {
                            seed      = seed * 1664525u + 1013904223u;
                            labels[i] = (seed >> 8) % nLabels;
                        }
                        if (labels.size() > 4) {
                            labels[2] = labels[1];                  // a repeated label
                            labels[3] = (labels[1] + M) % nLabels;  // and a sibling tied to the same mixture
                        }
                        Fsa::ConstAutomatonRef model = linearModel(labels, 3.0f, 0.0f, 30.0f);
                        Speech::Alignment      a, b, c;
                        double                 w0 = now();
                        refAligner.setModel(model, am);
                        refAligner.feed(segmentScorers(*gpuScaled, X, t0, t0 + T));
                        refAligner.getAlignment(a);
                        double w1 = now();
                        gpuAligner.setModel(model, am);
                        gpuAligner.feed(segmentScorers(*dfrScaled, X, t0, t0 + T));
                        gpuAligner.getAlignment(b);
                        double w2 = now();
                        secRef += w1 - w0;
                        secGpu += w2 - w1;
                        AlignDiff d = compareAlignments(a, b, mode ? 1e-4 : 0.0);
                        nItems += d.items;
                        nDiffer += d.differing + (refAligner.reachedFinalState() != gpuAligner.reachedFinalState());
                        nScoreBits += !d.scoreBits;
                        maxDiff = std::max(maxDiff, d.maxDiff);
                        maxRel  = std::max(maxRel, d.scoreRel);
                        nReached += gpuAligner.reachedFinalState();
                        if (kernel == 1) {  // direct kernel: bit-identical scores to the reference's CPU scorer
                            cpuAligner.setModel(model, am);
                            cpuAligner.feed(segmentScorers(*cpuScaled, X, t0, t0 + T));
                            cpuAligner.getAlignment(c);
                            AlignDiff e = compareAlignments(c, b, mode ? 1e-4 : 0.0);
                            nCpuDiffer += e.differing + (mode == 0 && !e.scoreBits);
                        }
                        t0 += T;
                        ++nSeg;
                    }
                    const MmGpu::GpuFeatureScorer* g  = dynamic_cast<const MmGpu::GpuFeatureScorer*>(dfr.get());
                    const bool                     ok = nSeg > 0 && nReached > 0 && nDiffer == 0 && nCpuDiffer == 0 &&
                                    (mode ? maxRel <= 1e-5 : nScoreBits == 0) && g->nBatches() == 0;
                    char name[96];
                    snprintf(name, sizeof name, "GpuAligner %s / %s kernel", mode ? "baum-welch" : "viterbi", kernel ? "direct" : "auto");
                    report(name, ok,
                           "segments=%zu reached=%zu items=%zu differing=%zu (vs CPU scorer: %zu) max|dw|=%.2g score %s; full "
                           "scoring passes=%u; s: reference %.3f, gpu %.3f",
                           nSeg, nReached, nItems, nDiffer, nCpuDiffer, maxDiff, mode ? (maxRel <= 1e-5 ? "rel<=1e-5" : "DIFFERS") : (nScoreBits ? "DIFFERS" : "bit-equal"),
                           g->nBatches(), secRef, secGpu);
                }
        }

        // ---- 10: frames/s of the plugin path against the C-ABI call it wraps -----------------------------------------
        const size_t timingFrames = args.count("timing-frames") ? atoi(args["timing-frames"].c_str()) : 0;
        if (timingFrames > 0) {
            Core::Configuration            gc  = scorerConfig("gpu", "gpu-diagonal-maximum", 0, 0);
            Core::Ref<Mm::FeatureScorer>   gpu = mm.createFeatureScorer(gc, ms);
            const MmGpu::GpuFeatureScorer* g   = dynamic_cast<const MmGpu::GpuFeatureScorer*>(gpu.get());
            const u32                      ld  = (M + 3) / 4 * 4;
            const int                      reps = 12;
            float* hx = static_cast<float*>(gmm_host_alloc(timingFrames * D * sizeof(float)));
            float* hs = static_cast<float*>(gmm_host_alloc(timingFrames * size_t(ld) * sizeof(float)));
            for (size_t t = 0; t < timingFrames; ++t)
                std::copy(X[t % X.size()].begin(), X[t % X.size()].end(), hx + t * D);
            double best[2] = {1e30, 1e30};
            double sink    = 0.0;
            for (int r = 0; r < reps; ++r) {  // the first repetitions warm the pinned pool and the device buffers
                double t0 = now();
                gpu->reset();
                for (size_t t = 0; t < timingFrames; ++t)
                    gpu->addFeature(X[t % X.size()]);
                while (!gpu->bufferEmpty())
                    sink += gpu->flush()->score(0);  // the first read sends the segment to the GPU
                double t1 = now();
                if (gmm_score_host(g->handle(), timingFrames, hx, hs, 0, ld) != GMM_OK)
                    sink = std::nan("");
                double t2 = now();
                best[0] = std::min(best[0], t1 - t0);
                best[1] = std::min(best[1], t2 - t1);
            }
            gmm_host_free(hx);
            gmm_host_free(hs);
            const double plugin = timingFrames / best[0], direct = timingFrames / best[1];
            printf("TIMING frames=%zu mixtures=%u plugin_frames_per_s=%.0f gmm_score_host_frames_per_s=%.0f ratio=%.3f\n",
                   timingFrames, M, plugin, direct, plugin / direct);
            report("plugin path timing", std::isfinite(sink) && plugin >= 0.8 * direct, "ratio=%.3f (segment mode, one read per frame)",
                   plugin / direct);

            // the alignment of one segment of `timingFrames` frames, wall clock from the first addFeature to the items:
            // (a) the reference's aligner on the GPU scorer (full score matrix over the host link, search on one core),
            // (b) MmGpu::GpuAligner on a deferred-scoring GPU scorer (compact block in HBM, search on the device)
            std::vector<Mm::FeatureVector> seg(timingFrames);
            for (size_t t = 0; t < timingFrames; ++t)
                seg[t] = X[t % X.size()];
            std::vector<u32> tie(M), labels(std::max<size_t>(1, timingFrames / 3));
            for (u32 l = 0; l < M; ++l)
                tie[l] = l;
            u32 seed = 99;
            for (size_t i = 0; i < labels.size(); ++i) {
                seed      = seed * 1664525u + 1013904223u;
                labels[i] = (seed >> 8) % M;
            }
            Core::Ref<const Am::AcousticModel> am(new TableAcousticModel(select("acoustic-model"), M, tie));
            Fsa::ConstAutomatonRef             model = linearModel(labels, 3.0f, 0.0f, 30.0f);
            Core::Configuration                ac(config, "aligner");
            ac.set(ac.getSelection() + ".min-acoustic-pruning", "1e30");
            ac.set(ac.getSelection() + ".max-acoustic-pruning", "1e30");
            Core::Configuration dc = scorerConfig("gpu-deferred", "gpu-diagonal-maximum", 0, 0);
            dc.set(dc.getSelection() + ".deferred-scoring", "true");
            Core::Ref<Mm::FeatureScorer> dfr = mm.createFeatureScorer(dc, ms);
            Search::Aligner              refAligner(ac);
            MmGpu::GpuAligner            gpuAligner(ac);
            double                       alignBest[2] = {1e30, 1e30};
            Speech::Alignment            a, b;
            for (int r = 0; r < 6; ++r) {
                double t0 = now();
                refAligner.setModel(model, am);
                refAligner.feed(segmentScorers(*gpu, seg, 0, timingFrames));
                refAligner.getAlignment(a);
                double t1 = now();
                gpuAligner.setModel(model, am);
                gpuAligner.feed(segmentScorers(*dfr, seg, 0, timingFrames));
                gpuAligner.getAlignment(b);
                double t2 = now();
                alignBest[0] = std::min(alignBest[0], t1 - t0);
                alignBest[1] = std::min(alignBest[1], t2 - t1);
            }
            AlignDiff d = compareAlignments(a, b, 0.0);
            printf("ALIGN-TIMING frames=%zu mixtures=%u labels=%zu reference_aligner_on_gpu_scorer_ms=%.3f gpu_aligner_ms=%.3f speedup=%.2f\n",
                   timingFrames, M, labels.size(), alignBest[0] * 1e3, alignBest[1] * 1e3, alignBest[0] / alignBest[1]);
            report("GpuAligner timing segment", d.differing == 0 && d.scoreBits && gpuAligner.reachedFinalState(),
                   "identical items=%zu; %.2fx the reference aligner fed by the same GPU scorer", d.items, alignBest[0] / alignBest[1]);
        }
        return failures_ ? 1 : 0;
    }
};

APPLICATION(CheckApp)
