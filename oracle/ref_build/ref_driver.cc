// TEST INFRASTRUCTURE (oracle/): command-line harness around the UNMODIFIED reference scorer
// classes of /root/reference/src/Mm.  Never linked into, imported by or executed from the product.
//
// It builds a Mm::MixtureSet from a raw "GMS1" model file (format: rasr_b200/gms.py), runs one of the
// reference scorers eagerly over all mixtures for every frame of a raw f32 feature file, and writes
//   <out>.score.f32  [T x M]   Scorer::score(m)                    (FeatureScorer.hh:39-40)
//   <out>.best.u32   [T x M]   AssigningScorer::bestDensity(m)     (AssigningFeatureScorer.hh:110-121)
// and, for scorer=simd, the tables the reference produced on the host (SimdFeatureScorer.cc:64-104):
//   <out>.q.i32 [T x M] quantizedScore().first, <out>.qmean.u8 [G x Dpad], <out>.qconst.i32 [G],
//   <out>.qcov.u32 [G], <out>.isrs.f32 [C x D] (scaled inverse std-dev), <out>.qx.u8 [T x C x Dpad],
//   <out>.meta.txt (scalingSquared, Dpad, G, C).
// Timing mode (threads=N repeat=R): one scorer instance per thread (the reference scorers are not
// thread-safe), frames split contiguously, wall clock of the slowest thread, best of R.
#include "ref_common.hh"

namespace {

double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

class RefScorerApp : public Core::Application {
    std::map<std::string, std::string> args_;

    std::string arg(const std::string& key, const std::string& dflt = "") const {
        std::map<std::string, std::string>::const_iterator i = args_.find(key);
        return i == args_.end() ? dflt : i->second;
    }

public:
    RefScorerApp() {
        setTitle("ref-scorer");
        setDefaultLoadConfigurationFile(false);
        setDefaultOutputXmlHeader(false);
    }
    std::string getUsage() const {
        return "ref_scorer model=<gms> features=<f32> scorer=max|sum|simd|batchint|batchfloat out=<prefix> "
               "[threads=N repeat=R dump=0|1 frames=T]\n";
    }

    // Eager evaluation through the assigning interface (AssigningFeatureScorer.hh:66-70,110-121).
    template<class S>
    void dumpAssigning(const S& fs, const std::vector<Mm::FeatureVector>& X, u32 M, const std::string& out) {
        std::vector<f32> score(X.size() * size_t(M));
        std::vector<u32> best(X.size() * size_t(M));
        for (size_t t = 0; t < X.size(); ++t) {
            Mm::AssigningFeatureScorer::AssigningScorer s = fs.getAssigningScorer(X[t]);
            for (u32 m = 0; m < M; ++m) {
                score[t * M + m] = s->score(m);
                best[t * M + m]  = s->bestDensity(m);
            }
        }
        writeArray(out + ".score.f32", score);
        writeArray(out + ".best.u32", best);
    }

    // Buffered protocol exactly as the callers drive it (Speech/AlignmentNode.cc:283-306).
    template<class S>
    void dumpBuffered(const S& fs, const std::vector<Mm::FeatureVector>& X, u32 M, const std::string& out) {
        std::vector<f32> score;
        score.reserve(X.size() * size_t(M));
        fs.reset();
        for (size_t t = 0; t < X.size(); ++t) {
            if (fs.isBuffered() && !fs.bufferFilled()) {
                fs.addFeature(X[t]);
            }
            else {
                Mm::FeatureScorer::Scorer s = fs.getScorer(X[t]);
                for (u32 m = 0; m < M; ++m)
                    score.push_back(s->score(m));
            }
        }
        while (!fs.bufferEmpty()) {
            Mm::FeatureScorer::Scorer s = fs.flush();
            for (u32 m = 0; m < M; ++m)
                score.push_back(s->score(m));
        }
        writeArray(out + ".score.f32", score);
    }

    // Tables and integer results of the SIMD scorer (SimdFeatureScorer.cc:64-104,158-176).
    void dumpSimd(const Mm::SimdGaussDiagonalMaximumFeatureScorer& fs, const std::vector<Mm::FeatureVector>& X,
                  u32 M, u32 D, const std::string& out) {
        typedef Mm::SimdGaussDiagonalMaximumFeatureScorer S;
        const size_t     C = fs.covarianceTable_.size();
        std::vector<f32> isrs;
        for (size_t c = 0; c < C; ++c)
            isrs.insert(isrs.end(), fs.covarianceTable_[c].inverseSquareRootDiagonal().begin(),
                        fs.covarianceTable_[c].inverseSquareRootDiagonal().end());
        std::vector<u8>  qmean;
        std::vector<s32> qconst;
        std::vector<u32> qcov;
        size_t           Dpad = 0;
        for (u32 m = 0; m < M; ++m) {
            const S::MixtureElement& me = fs.mixtureTable_[m];
            for (size_t k = 0; k < me.nDensities(); ++k) {
                const S::MixtureElement::Density& d = me.density(k);
                Dpad                                = d.preparedMean_.size();
                qmean.insert(qmean.end(), d.preparedMean_.begin(), d.preparedMean_.end());
                qconst.push_back(d.constantWeight_);
                qcov.push_back(d.covarianceIndex_);
            }
        }
        std::vector<s32> q(X.size() * size_t(M));
        std::vector<u8>  qx;
        for (size_t t = 0; t < X.size(); ++t) {
            std::vector<S::PreparedFeatureVector> fpc = fs.multiplyAndQuantize(X[t]);
            for (size_t c = 0; c < C; ++c)
                qx.insert(qx.end(), fpc[c].begin(), fpc[c].end());
            for (u32 m = 0; m < M; ++m)
                q[t * M + m] = fs.quantizedScore(fs.mixtureTable_[m], fpc).first;
        }
        writeArray(out + ".q.i32", q);
        writeArray(out + ".qmean.u8", qmean);
        writeArray(out + ".qconst.i32", qconst);
        writeArray(out + ".qcov.u32", qcov);
        writeArray(out + ".isrs.f32", isrs);
        writeArray(out + ".qx.u8", qx);
        FILE* f = fopen((out + ".meta.txt").c_str(), "w");
        fprintf(f, "scalingSquared %.9g\nDpad %zu\nG %zu\nC %zu\nD %u\n", (double)fs.scalingSquared_, Dpad,
                qconst.size(), C, D);
        fclose(f);
    }

    // One scorer per thread, frames split contiguously; returns best-of-R seconds for all T frames.
    template<class S, bool buffered>
    double timeScorer(Core::Ref<const Mm::MixtureSet> ms, const std::string& sel,
                      const std::vector<Mm::FeatureVector>& X, u32 M, int nThreads, int repeat, double& checksum) {
        std::vector<S*> scorers;
        for (int i = 0; i < nThreads; ++i)
            scorers.push_back(new S(select(sel), ms));
        std::vector<double> sums(nThreads, 0.0);
        double              best = 1e300;
        for (int r = 0; r < repeat; ++r) {
            std::vector<std::thread> pool;
            std::vector<double>      secs(nThreads, 0.0);
            for (int i = 0; i < nThreads; ++i) {
                size_t lo = X.size() * i / nThreads, hi = X.size() * (i + 1) / nThreads;
                pool.push_back(std::thread([&, i, lo, hi]() {
                    const S& fs  = *scorers[i];
                    double   acc = 0.0, t0 = now();
                    if (!buffered) {
                        for (size_t t = lo; t < hi; ++t) {
                            Mm::FeatureScorer::Scorer s = fs.getScorer(X[t]);
                            for (u32 m = 0; m < M; ++m)
                                acc += s->score(m);
                        }
                    }
                    else {
                        fs.reset();
                        for (size_t t = lo; t < hi; ++t) {
                            if (!fs.bufferFilled()) {
                                fs.addFeature(X[t]);
                                continue;
                            }
                            Mm::FeatureScorer::Scorer s = fs.getScorer(X[t]);
                            for (u32 m = 0; m < M; ++m)
                                acc += s->score(m);
                        }
                        while (!fs.bufferEmpty()) {
                            Mm::FeatureScorer::Scorer s = fs.flush();
                            for (u32 m = 0; m < M; ++m)
                                acc += s->score(m);
                        }
                    }
                    secs[i] = now() - t0;
                    sums[i] = acc;
                }));
            }
            for (size_t i = 0; i < pool.size(); ++i)
                pool[i].join();
            best = std::min(best, *std::max_element(secs.begin(), secs.end()));
        }
        checksum = 0.0;
        for (int i = 0; i < nThreads; ++i) {
            checksum += sums[i];
            delete scorers[i];
        }
        return best;
    }

    int main(const std::vector<std::string>& arguments) {
        for (size_t i = 0; i < arguments.size(); ++i) {
            size_t eq = arguments[i].find('=');
            if (eq != std::string::npos)
                args_[arguments[i].substr(0, eq)] = arguments[i].substr(eq + 1);
        }
        const std::string kind = arg("scorer", "max"), out = arg("out", "ref_out");
        RawModel          raw;
        if (!loadRawModel(arg("model"), raw)) {
            std::cerr << "cannot read model '" << arg("model") << "'\n" << getUsage();
            return 2;
        }
        std::vector<f32> flat;
        {
            std::ifstream in(arg("features").c_str(), std::ios::binary | std::ios::ate);
            if (!in) {
                std::cerr << "cannot read features\n";
                return 2;
            }
            size_t bytes = in.tellg();
            in.seekg(0);
            flat.resize(bytes / 4);
            in.read(reinterpret_cast<char*>(flat.data()), flat.size() * 4);
        }
        size_t T = flat.size() / raw.D;
        if (!arg("frames").empty())
            T = std::min(T, (size_t)atol(arg("frames").c_str()));
        std::vector<Mm::FeatureVector> X(T, Mm::FeatureVector(raw.D));
        for (size_t t = 0; t < T; ++t)
            std::copy(flat.begin() + t * raw.D, flat.begin() + (t + 1) * raw.D, X[t].begin());

        Core::Ref<Mm::MixtureSet>       msw = buildMixtureSet(raw);
        Core::Ref<const Mm::MixtureSet> ms(msw);
        const u32                       M = raw.M;

        if (arg("dump", "1") != "0") {
            if (kind == "max") {
                Mm::GaussDiagonalMaximumFeatureScorer fs(select("fs"), ms);
                dumpAssigning(fs, X, M, out);
            }
            else if (kind == "sum") {
                Mm::GaussDiagonalSumFeatureScorer fs(select("fs"), ms);
                dumpAssigning(fs, X, M, out);
            }
            else if (kind == "simd") {
                Mm::SimdGaussDiagonalMaximumFeatureScorer fs(select("fs"), ms);
                dumpAssigning(fs, X, M, out);
                dumpSimd(fs, X, M, raw.D, out);
            }
            else if (kind == "batchint") {
                Mm::BatchIntFeatureScorer fs(select("fs"), ms);
                dumpBuffered(fs, X, M, out);
            }
            else if (kind == "batchfloat") {
                Mm::BatchFloatFeatureScorer fs(select("fs"), ms);
                dumpBuffered(fs, X, M, out);
            }
            else {
                std::cerr << getUsage();
                return 2;
            }
        }
        int threads = atoi(arg("threads", "0").c_str()), repeat = std::max(1, atoi(arg("repeat", "1").c_str()));
        if (threads > 0) {
            double secs = 0.0, checksum = 0.0;
            if (kind == "max")
                secs = timeScorer<Mm::GaussDiagonalMaximumFeatureScorer, false>(ms, "fs", X, M, threads, repeat, checksum);
            else if (kind == "sum")
                secs = timeScorer<Mm::GaussDiagonalSumFeatureScorer, false>(ms, "fs", X, M, threads, repeat, checksum);
            else if (kind == "simd")
                secs = timeScorer<Mm::SimdGaussDiagonalMaximumFeatureScorer, false>(ms, "fs", X, M, threads, repeat, checksum);
            else if (kind == "batchint")
                secs = timeScorer<Mm::BatchIntFeatureScorer, true>(ms, "fs", X, M, threads, repeat, checksum);
            else if (kind == "batchfloat")
                secs = timeScorer<Mm::BatchFloatFeatureScorer, true>(ms, "fs", X, M, threads, repeat, checksum);
            FILE* f = fopen((out + ".timing.json").c_str(), "w");
            fprintf(f, "{\"scorer\": \"%s\", \"frames\": %zu, \"mixtures\": %u, \"threads\": %d, \"repeat\": %d, "
                       "\"seconds\": %.6f, \"frames_per_s\": %.3f, \"checksum\": %.6e}\n",
                    kind.c_str(), T, M, threads, repeat, secs, T / secs, checksum);
            fclose(f);
        }
        return 0;
    }
};

APPLICATION(RefScorerApp)
