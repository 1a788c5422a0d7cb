// TEST INFRASTRUCTURE (oracle/): drives the reference's own, UNMODIFIED Search::Aligner (src/Search/Aligner.cc) with
// its Fsa library (Fsa::posterior64, Fsa::transpose, ...) and Speech::Alignment (combineItems, clipWeights,
// shiftMinToZeroWeights, expm, normalizeWeights), compiled by oracle/ref_build/Makefile with the reference's flags, to
// pin oracle/gmm_oracle.c orc_viterbi_align / orc_baum_welch_align and the fixtures tests/golden/align_*.npz.
//
// What is NOT the reference's here, and why it does not matter for the pin:
//   * the acoustic model is a test double (TableAcousticModel): the aligner asks it for emissionIndex(label) only
//     (Aligner.cc:225) and for the alphabet it attaches to the result (Aligner.cc:718); labels ARE emission indices;
//   * the per-frame scorers are a test double (RowScorer) returning one row of a score block -- the block is what the
//     scoring path hands the aligner, computed elsewhere (ref_scorer pins that part);
//   * libxml2 is absent: stubs/libxml/parser.h holds the handful of type names Core/XmlParser.hh mentions so that the
//     HEADERS parse; no XML parsing TU is compiled, stubs/align_stubs.cc aborts in the few word-lattice symbols
//     (Aligner::WordLatticeBuilder) the aligner's object file refers to and this driver never reaches.
//
// usage: ref_aligner in=<file> out=<prefix> mode=viterbi|baum-welch [labels=<file>]
//            --*.aligner.min-acoustic-pruning=1e30 --*.aligner.max-acoustic-pruning=1e30   (unpruned, one pass)
//   in : u32 T, M, S, A, initial; u32 arcOffsets[S+1]; u32 arcTarget[A]; u32 arcEmission[A]; f32 arcWeight[A];
//        f32 finalWeight[S] (+inf: not final); f32 scores[T*M]
//   labels (optional): u32 nLabels; u32 emissionOfLabel[nLabels] -- the arcs' "emission" field is then a LABEL (an
//        allophone state in RASR) the acoustic model ties to an emission; items name the label, as in RASR
//   out: <prefix>.items.bin = u32 n, n x {u32 time, u32 emission, f64 weight} in the order Alignment holds them;
//        <prefix>.score.f32 = {alignmentScore(), total of the posterior automaton (baum-welch) or the same (viterbi)}
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "ref_common.hh"  // Mm::FeatureScorer::ContextScorer is protected: the harness switch makes it reachable

#include <Am/AcousticModel.hh>
#include <Core/Application.hh>
#include <Fsa/Static.hh>
#include <Mm/FeatureScorer.hh>
#include <Search/Aligner.hh>
#include <Speech/Alignment.hh>

#include "ref_table_am.hh"

namespace {

class RowScorer : public Mm::FeatureScorer::ContextScorer {
    const f32* row_;
    u32        n_;

public:
    RowScorer(const f32* row, u32 n)
            : row_(row), n_(n) {}
    virtual Mm::EmissionIndex nEmissions() const {
        return n_;
    }
    virtual Mm::Score score(Mm::EmissionIndex e) const {
        return row_[e];
    }
};

}  // namespace

class RefAlignerApp : public Core::Application {
    std::map<std::string, std::string> args_;

    std::string arg(const std::string& key, const std::string& dflt = "") const {
        std::map<std::string, std::string>::const_iterator i = args_.find(key);
        return i == args_.end() ? dflt : i->second;
    }

public:
    RefAlignerApp() {
        setTitle("ref-aligner");
        setDefaultLoadConfigurationFile(false);
        setDefaultOutputXmlHeader(false);
    }
    std::string getUsage() const {
        return "ref_aligner in=<file> out=<prefix> mode=viterbi|baum-welch\n";
    }

    int main(const std::vector<std::string>& arguments) {
        for (size_t i = 0; i < arguments.size(); ++i) {
            size_t eq = arguments[i].find('=');
            if (eq != std::string::npos)
                args_[arguments[i].substr(0, eq)] = arguments[i].substr(eq + 1);
        }
        const std::string in = arg("in"), out = arg("out"), mode = arg("mode", "viterbi");
        FILE*             f = fopen(in.c_str(), "rb");
        if (!f || out.empty()) {
            std::cerr << getUsage();
            return 2;
        }
        u32 hdr[5];
        if (fread(hdr, 4, 5, f) != 5)
            return 3;
        const u32        T = hdr[0], M = hdr[1], S = hdr[2], A = hdr[3], initial = hdr[4];
        std::vector<u32> arcOffsets, arcTarget, arcEmission;
        std::vector<f32> arcWeight, finalWeight, scores;
        if (!readArray(f, arcOffsets, S + 1) || !readArray(f, arcTarget, A) || !readArray(f, arcEmission, A) ||
            !readArray(f, arcWeight, A) || !readArray(f, finalWeight, S) || !readArray(f, scores, size_t(T) * M))
            return 4;
        fclose(f);

        // the model acceptor: states in id order, arcs in the order given (= the order the aligner expands them)
        Fsa::StaticAutomaton* model = new Fsa::StaticAutomaton();
        model->setType(Fsa::TypeAcceptor);
        model->setSemiring(Fsa::TropicalSemiring);
        for (u32 s = 0; s < S; ++s) {
            Fsa::State* sp = new Fsa::State(s);
            u32 bits;  // -ffast-math folds isinf() away: test the bit pattern of +inf
            memcpy(&bits, &finalWeight[s], 4);
            if (bits != 0x7f800000u)
                sp->setFinal(Fsa::Weight(finalWeight[s]));
            for (u32 a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a)
                sp->newArc(arcTarget[a], Fsa::Weight(arcWeight[a]), arcEmission[a]);
            model->setState(sp);
        }
        model->setInitialStateId(initial);
        Fsa::ConstAutomatonRef modelRef(model);

        // unpruned: one pass at a threshold nothing exceeds (Aligner.cc:584-611); min-weight 0 keeps every item
        // (the caller passes --*.aligner.min-acoustic-pruning=1e30 --*.aligner.max-acoustic-pruning=1e30)
        std::vector<u32> emissionOfLabel;
        if (!arg("labels").empty()) {  // u32 nLabels; u32 emissionOfLabel[nLabels]: arcs then carry labels
            FILE* lf = fopen(arg("labels").c_str(), "rb");
            u32   nLabels = 0;
            if (!lf || fread(&nLabels, 4, 1, lf) != 1 || !readArray(lf, emissionOfLabel, nLabels))
                return 4;
            fclose(lf);
        }
        Core::Ref<const Am::AcousticModel> am(new TableAcousticModel(select("acoustic-model"), M, emissionOfLabel));
        Search::Aligner                    aligner(select("aligner"));
        aligner.selectMode(mode == "viterbi" ? Search::Aligner::modeViterbi : Search::Aligner::modeBaumWelch);
        aligner.setModel(modelRef, am);
        std::vector<Mm::FeatureScorer::Scorer> scorers;
        for (u32 t = 0; t < T; ++t)
            scorers.push_back(Mm::FeatureScorer::Scorer(new RowScorer(scores.data() + size_t(t) * M, M)));
        aligner.feed(scorers);

        f32               result[2] = {aligner.alignmentScore(), aligner.alignmentScore()};
        Speech::Alignment alignment;
        if (aligner.reachedFinalState()) {
            std::pair<Fsa::ConstAutomatonRef, Fsa::Weight> posterior =
                    aligner.getAlignmentPosteriorFsa(aligner.getAlignmentFsa());
            if (mode != "viterbi")
                result[1] = f32(posterior.second);
            aligner.getAlignment(alignment, posterior);
        }
        f = fopen((out + ".items.bin").c_str(), "wb");
        if (!f)
            return 5;
        const u32 n = alignment.size();
        fwrite(&n, 4, 1, f);
        for (Speech::Alignment::const_iterator it = alignment.begin(); it != alignment.end(); ++it) {
            const u32 te[2] = {u32(it->time), u32(it->emission)};
            const f64 w     = it->weight;
            fwrite(te, 4, 2, f);
            fwrite(&w, 8, 1, f);
        }
        fclose(f);
        f = fopen((out + ".score.f32").c_str(), "wb");
        if (!f)
            return 5;
        fwrite(result, 4, 2, f);
        fclose(f);
        return 0;
    }
};

APPLICATION(RefAlignerApp)
