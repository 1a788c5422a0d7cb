// TEST INFRASTRUCTURE (oracle/): command-line harness around the UNMODIFIED reference estimator stack and text model
// reader / writer of /root/reference/src/Mm.  Never linked into, imported by or executed from the product.
// It pins SURVEY.md 8(f) ranks 2 and 3 (model files, training E-step / M-step):
//
//   ref_estimator model=<gms> out=<prefix> [features=<f32> align=<u32 mixture per frame> weights=<f64 per frame>
//                 mode=viterbi|baum-welch scorer=max|sum min-variance=<v> threshold=<w>]
//
//   <out>.pms          MixtureSet::write(std::ostream&)                       (Mm/MixtureSet.cc:142-169)
//   <out>.reread.pms   the same file read back by MixtureSet::read and written again   (:170-214)
// and, with features:
//   <out>.best.u32     density chosen for every aligned frame (viterbi), AbstractMixtureSetEstimator::densityIndex
//   <out>.mixset       the accumulators, MixtureSetEstimator::write(BinaryOutputStream&)
//                      (Mm/AbstractMixtureSetEstimator.cc:482-506) after accumulate(mixture, feature[, weight]) of every
//                      frame (:117-147): viterbi through scorer->bestDensity, baum-welch through
//                      getDensityPosteriorProbabilities (:386-395) with the weight threshold
//   <out>.mixset2      the same accumulators read back (read(), :433-480) and written again
//   <out>.est.pms      estimate() (:299-335) of those accumulators, written as text
#define REF_WITH_ESTIMATORS
#include "ref_common.hh"

class RefEstimatorApp : public Core::Application {
    std::map<std::string, std::string> args_;

    std::string arg(const std::string& key, const std::string& dflt = "") const {
        std::map<std::string, std::string>::const_iterator i = args_.find(key);
        return i == args_.end() ? dflt : i->second;
    }

public:
    RefEstimatorApp() {
        setTitle("ref-estimator");
        setDefaultLoadConfigurationFile(false);
        setDefaultOutputXmlHeader(false);
    }
    std::string getUsage() const {
        return "ref_estimator model=<gms> out=<prefix> [features=<f32> align=<u32> weights=<f64> mode=viterbi|baum-welch "
               "scorer=max|sum min-variance=<v> threshold=<w>]\n";
    }

    template<class T>
    bool readFile(const std::string& path, std::vector<T>& v) {
        std::ifstream in(path.c_str(), std::ios::binary | std::ios::ate);
        if (!in)
            return false;
        size_t bytes = in.tellg();
        in.seekg(0);
        v.resize(bytes / sizeof(T));
        in.read(reinterpret_cast<char*>(v.data()), v.size() * sizeof(T));
        return true;
    }

    int main(const std::vector<std::string>& arguments) {
        for (size_t i = 0; i < arguments.size(); ++i) {
            size_t eq = arguments[i].find('=');
            if (eq != std::string::npos)
                args_[arguments[i].substr(0, eq)] = arguments[i].substr(eq + 1);
        }
        const std::string out = arg("out", "ref_est");
        RawModel          raw;
        if (!loadRawModel(arg("model"), raw)) {
            std::cerr << "cannot read model\n" << getUsage();
            return 2;
        }
        Core::Ref<Mm::MixtureSet>       msw = buildMixtureSet(raw);
        Core::Ref<const Mm::MixtureSet> ms(msw);

        // ---- text model format: write, read back, write again ----
        {
            std::ofstream o((out + ".pms").c_str());
            if (!ms->write(o))
                return 3;
        }
        {
            Mm::MixtureSet again;
            std::ifstream  i((out + ".pms").c_str());
            again.read(i);
            std::ofstream o((out + ".reread.pms").c_str());
            if (!again.write(o))
                return 3;
        }
        if (arg("features").empty())
            return 0;

        // ---- accumulation ----
        std::vector<f32> flat;
        std::vector<u32> align;
        std::vector<f64> weights;
        if (!readFile(arg("features"), flat) || !readFile(arg("align"), align)) {
            std::cerr << "cannot read features / alignment\n";
            return 2;
        }
        const bool weighted = !arg("weights").empty();
        if (weighted && !readFile(arg("weights"), weights))
            return 2;
        const size_t T = align.size();
        if (flat.size() != T * raw.D || (weighted && weights.size() != T)) {
            std::cerr << "features / alignment / weights disagree\n";
            return 2;
        }
        Core::Ref<const Mm::AssigningFeatureScorer> scorer;
        const Mm::GaussDiagonalSumFeatureScorer* sumScorer = 0;
        if (arg("scorer", "max") == "sum")
            scorer = Core::ref(sumScorer = new Mm::GaussDiagonalSumFeatureScorer(select("fs"), ms));
        else
            scorer = Core::ref(new Mm::GaussDiagonalMaximumFeatureScorer(select("fs"), ms));

        Mm::MixtureSetEstimator      est(select("est"));
        Mm::AbstractMixtureSetEstimator& acc = est;  // accumulate(mixture, feature[, weight]) is hidden in the subclass
        est.setTopology(ms);
        est.setAssigningFeatureScorer(scorer);
        est.viterbi_ = arg("mode", "viterbi") != "baum-welch";   // paramMode (AbstractMixtureSetEstimator.cc:45-49)
        if (!arg("threshold").empty())
            est.setWeightThreshold(atof(arg("threshold").c_str()));
        if (!arg("min-variance").empty())
            est.minVariance_ = (Mm::VarianceType)atof(arg("min-variance").c_str());

        std::vector<u32> best(T, 0xFFFFFFFFu);
        for (size_t t = 0; t < T; ++t) {
            if (align[t] == 0xFFFFFFFFu)
                continue;
            Core::Ref<const Mm::Feature::Vector> x(new Mm::Feature::Vector(
                    Mm::FeatureVector(flat.begin() + t * raw.D, flat.begin() + (t + 1) * raw.D)));
            if (sumScorer)
                sumScorer->lastContext_ = 0;  // see below
            if (est.viterbi_) {
                Core::Ref<const Mm::Feature> f(new Mm::Feature(x));
                best[t] = scorer->getAssigningScorer(f)->bestDensity(align[t]);
            }
            // GaussDiagonalSumFeatureScorer caches the density scores of (context ADDRESS, mixture)
            // (GaussDiagonalMaximumFeatureScorer.cc:239-246; both members are never initialised, :222-228).  The
            // estimator frees its context after every frame (AbstractMixtureSetEstimator.cc:386-395), the next one is
            // allocated at the same address, and two consecutive frames aligned to the same mixture then reuse the
            // PREVIOUS frame's density scores.  That artefact of the allocator is not part of the algorithm: the
            // harness forgets the cached address before every frame.
            if (sumScorer)
                sumScorer->lastContext_ = 0;
            if (weighted)
                acc.accumulate(align[t], x, weights[t]);
            else
                acc.accumulate(align[t], x);
        }
        writeArray(out + ".best.u32", best);
        {
            Core::BinaryOutputStream bos(out + ".mixset");
            est.write(bos);
        }
        {
            Mm::MixtureSetEstimator again(select("est"));
            Core::BinaryInputStream bis(out + ".mixset");
            again.read(bis);
            Core::BinaryOutputStream bos(out + ".mixset2");
            again.write(bos);
        }
        // ---- M-step ----
        {
            Core::Ref<Mm::MixtureSet> estimated = est.estimate();
            std::ofstream             o((out + ".est.pms").c_str());
            if (!estimated->write(o))
                return 3;
        }
        return 0;
    }
};

APPLICATION(RefEstimatorApp)
