/** RASR-side plugin, second part: the alignment consumer behind the interface of Search::Aligner.
 *
 *  Search::Aligner (src/Search/Aligner.hh:36-176) is what Speech::AlignmentNode drives per segment
 *  (src/Speech/AlignmentNode.cc:193, 211, 313-322): setModel(automaton, acoustic model), feed(scorers of every frame),
 *  reachedFinalState(), alignmentScore(), getAlignment(Speech::Alignment&).  MmGpu::GpuAligner offers those calls with
 *  the same meaning, for the scorers of a GpuFeatureScorer: the frames the scorers stand for go to the device ONCE,
 *  only the emissions the automaton names are scored (the reference's lazy score(e)), the score block stays in HBM and
 *  the search runs there (include/gmm_b200.h gmm_score_align_labels_host / gmm_score_baum_welch_labels_host); what
 *  returns is the alignment -- 4 bytes per frame in Viterbi mode, one posterior row per frame in Baum-Welch mode.
 *
 *  Same results as the reference's aligner run UNPRUNED (its one pass at max-acoustic-pruning when that is large enough
 *  to keep every hypothesis; Aligner.cc:584-611) with min-weight 0: Viterbi items and score bit for bit, exact f32 ties
 *  resolved in the reference's hypothesis order; Baum-Welch weights within 1e-4.  rasr/check.cc runs both side by side.
 *
 *  Not offered: pruned search, getAlignmentFsa() / word lattices (the path leaves the device as items, not as an Fsa),
 *  use-partial-sums.  mode = baum-welch with min-weight > 0 is refused (criticalError), never silently approximated.
 */
#ifndef _MM_GPU_ALIGNER_HH
#define _MM_GPU_ALIGNER_HH

#include <vector>

#include <Am/AcousticModel.hh>
#include <Core/Component.hh>
#include <Fsa/Automaton.hh>
#include <Mm/FeatureScorer.hh>
#include <Speech/Alignment.hh>

#include "GpuFeatureScorer.hh"

namespace MmGpu {

class GpuAligner : public Core::Component {
public:
    typedef Speech::Score Score;
    enum Mode { modeViterbi, modeBaumWelch };  // Search::Aligner::Mode

    static const Core::Choice          choiceMode;
    static const Core::ParameterChoice paramMode;       // "mode", as Aligner.cc:498-499
    static const Core::ParameterFloat  paramMinWeight;  // "min-weight", as Aligner.cc:501-504: must stay 0

    GpuAligner(const Core::Configuration&);
    ~GpuAligner();

    /** Flattens the automaton: states in id order, arcs in Fsa order (the order the reference expands them in), every
     *  distinct arc label a slot whose mixture is acousticModel->emissionIndex(label) (Aligner.cc:225). */
    void setModel(Fsa::ConstAutomatonRef model, Core::Ref<const Am::AcousticModel>);
    /** as Search::Aligner::getModel (AlignmentNode.cc:193 reuses the model of the previous segment through it) */
    Fsa::ConstAutomatonRef getModel() const {
        return model_;
    }
    void restart();

    /** The scorers must come from a GpuFeatureScorer on the float engine (directly, or through the scaling wrapper
     *  Mm::FeatureScorerScaling, whose scale is then applied on the device): criticalError otherwise -- there is no
     *  host path.  Scorers whose frames were never asked for a score cost no scoring pass of their own. */
    void feed(const std::vector<Mm::FeatureScorer::Scorer>&);

    Score alignmentScore() const {
        return score_;
    }
    bool reachedFinalState() const {
        return score_ != Core::Type<Score>::max;
    }
    /** Items as Aligner::getAlignment leaves them: (time, allophone-state label, weight), sorted by time and, within a
     *  frame, by decreasing weight (Alignment::sortItems(false) on -log weights, Aligner.cc:697-700); weight 1 in
     *  Viterbi mode, the normalised state posterior (> 0) in Baum-Welch mode; score and alphabet set. */
    void getAlignment(Speech::Alignment&) const;

    void selectMode(Mode mode) {
        viterbi_ = mode == modeViterbi;
    }
    Mode getMode() const {
        return viterbi_ ? modeViterbi : modeBaumWelch;
    }
    /** device calls so far (one per feed) */
    u32 nCalls() const {
        return nCalls_;
    }

private:
    Fsa::ConstAutomatonRef             model_;
    Core::Ref<const Am::AcousticModel> acousticModel_;
    bool                               viterbi_;
    // the flattened automaton (include/gmm_b200.h: one segment)
    std::vector<u32>          arcOffsets_, arcTarget_, arcLabelSlot_, labelMixture_;
    std::vector<Fsa::LabelId> labelOfSlot_;
    std::vector<f32>          arcWeight_, finalWeight_;
    u32                       initial_;
    bool                      haveModel_, initialIsFinal_;
    // the last feed
    u32              nFrames_;
    Score            score_;
    std::vector<u32> pathSlot_;  // Viterbi
    std::vector<f32> weights_;   // Baum-Welch: [nFrames x nLabels]
    u32              nCalls_;
};

}  // namespace MmGpu

#endif  // _MM_GPU_ALIGNER_HH
