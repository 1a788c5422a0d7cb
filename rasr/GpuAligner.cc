#include "GpuAligner.hh"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>

#include <Fsa/Static.hh>
#include <Mm/ScaledFeatureScorer.hh>

#include <gmm_b200.h>

using namespace MmGpu;

const Core::Choice GpuAligner::choiceMode(
        "viterbi", modeViterbi,
        "baum-welch", modeBaumWelch,
        Core::Choice::endMark());

const Core::ParameterChoice GpuAligner::paramMode(
        "mode", &choiceMode, "describes how alignments are created and observations are weighted", modeViterbi);

const Core::ParameterFloat GpuAligner::paramMinWeight(
        "min-weight",
        "alignment items with a weight (i.e. state posterior probability) below this threshold are pruned: "
        "the device aligner keeps every item, only 0 is accepted",
        0.0);

GpuAligner::GpuAligner(const Core::Configuration& c)
        : Core::Component(c),
          viterbi_(paramMode(c) == modeViterbi),
          initial_(0),
          haveModel_(false),
          nFrames_(0),
          score_(Core::Type<Score>::max),
          nCalls_(0) {
    // Aligner::getLogWeightThreshold (Aligner.cc:564-571): a min-weight above delta prunes the posterior lattice
    // (Fsa::prunePosterior re-weights it) -- not done on the device, so not accepted here
    if (paramMinWeight(c) > Core::Type<Mm::Weight>::delta)
        criticalError("the GPU aligner keeps every alignment item: min-weight must be 0");
}

GpuAligner::~GpuAligner() {}

void GpuAligner::setModel(Fsa::ConstAutomatonRef model, Core::Ref<const Am::AcousticModel> acousticModel) {
    require(model);
    require(model->initialStateId() != Fsa::InvalidStateId);  // as Aligner.cc:575-576
    model_         = model;
    acousticModel_ = acousticModel;
    Core::Ref<Fsa::StaticAutomaton> s = Fsa::staticCopy(model);  // what SearchSpace::setModel works on (Aligner.cc:117)
    const u32                       S = s->size();
    arcOffsets_.assign(1, 0);
    arcTarget_.clear();
    arcLabelSlot_.clear();
    arcWeight_.clear();
    labelMixture_.clear();
    labelOfSlot_.clear();
    // +inf marks a non-final state for the C-ABI.  RASR is built with -ffast-math, under which comparisons with an
    // infinity may be folded away: the value is only ever stored (bit pattern), finality is kept in initialIsFinal_
    f32       notFinal;
    const u32 infBits = 0x7f800000u;
    memcpy(&notFinal, &infBits, 4);
    finalWeight_.assign(S, notFinal);
    initialIsFinal_ = false;
    std::map<Fsa::LabelId, u32> slotOfLabel;
    for (Fsa::StateId id = 0; id < S; ++id) {
        const Fsa::State* st = s->fastState(id);
        if (st) {
            for (Fsa::State::const_iterator a = st->begin(); a != st->end(); ++a) {  // Fsa arc order = expansion order
                std::map<Fsa::LabelId, u32>::const_iterator slot = slotOfLabel.find(a->input());
                if (slot == slotOfLabel.end()) {
                    slot = slotOfLabel.insert(std::make_pair(a->input(), u32(labelOfSlot_.size()))).first;
                    labelOfSlot_.push_back(a->input());
                    labelMixture_.push_back(acousticModel_->emissionIndex(a->input()));  // Aligner.cc:225
                }
                arcTarget_.push_back(a->target());
                arcLabelSlot_.push_back(slot->second);
                arcWeight_.push_back(f32(a->weight()));
            }
            if (st->isFinal()) {
                finalWeight_[id] = f32(st->weight_);
                if (id == s->initialStateId())
                    initialIsFinal_ = true;
            }
        }
        arcOffsets_.push_back(arcTarget_.size());
    }
    initial_   = s->initialStateId();
    haveModel_ = true;
    restart();
}

void GpuAligner::restart() {
    nFrames_ = 0;
    score_   = Core::Type<Score>::max;
    pathSlot_.clear();
    weights_.clear();
}

void GpuAligner::feed(const std::vector<Mm::FeatureScorer::Scorer>& scorers) {
    require(haveModel_);
    restart();
    const u32 T = scorers.size();
    // the frames behind the scorers: the GPU scorer's contexts, possibly inside the scaling wrapper
    const GpuFeatureScorer*                       parent = 0;
    Mm::Score                                     scale  = 1;
    std::vector<const GpuFeatureScorer::Context*> ctx(T);
    for (u32 t = 0; t < T; ++t) {
        Mm::FeatureScorer::Scorer                                     inner = scorers[t];
        Mm::Score                                                     s     = 1;
        const Mm::FeatureScorerScaling::ScaledContextScorer* scaled =
                dynamic_cast<const Mm::FeatureScorerScaling::ScaledContextScorer*>(inner.get());
        if (scaled) {
            s     = scaled->getScale();
            inner = scaled->getUnscaledScorer();
        }
        ctx[t] = dynamic_cast<const GpuFeatureScorer::Context*>(inner.get());
        if (!ctx[t])
            criticalError("the GPU aligner aligns the frames of a GPU feature scorer: frame %d has another scorer", int(t));
        if (t == 0) {
            parent = ctx[t]->parent_;
            scale  = s;
        }
        else if (ctx[t]->parent_ != parent || s != scale)
            criticalError("the GPU aligner needs all frames of a segment from one feature scorer with one scale");
    }
    if (T && parent->quantized())
        criticalError("the GPU aligner runs on the float engine (gpu-diagonal-maximum / gpu-batch-diagonal-maximum-float)");
    if (T == 0 && !parent) {
        // no frame, no scorer to take an engine from: the start hypothesis is the only one (Aligner.cc:376-396)
        if (initialIsFinal_)
            score_ = finalWeight_[initial_];
        return;
    }
    const u32 D = parent->dimension();
    // frames of one segment normally lie in one page-locked block, one after the other: handed over in place
    const float*       x = 0;
    std::vector<float> gathered;
    bool               inPlace = true;
    for (u32 t = 1; t < T && inPlace; ++t)
        inPlace = ctx[t]->block_.get() == ctx[0]->block_.get() && ctx[t]->row_ == ctx[0]->row_ + t;
    if (T && inPlace)
        x = ctx[0]->block_->features + size_t(ctx[0]->row_) * D;
    else if (T) {
        gathered.resize(size_t(T) * D);
        for (u32 t = 0; t < T; ++t)
            std::copy(ctx[t]->block_->features + size_t(ctx[t]->row_) * D,
                      ctx[t]->block_->features + size_t(ctx[t]->row_ + 1) * D, gathered.begin() + size_t(t) * D);
        x = &gathered[0];
    }
    gmm_handle* h = parent->handle();
    parent->check(gmm_sync(h), "gmm_sync");  // a batch of the scorer itself may still be in flight on this engine
    parent->check(gmm_set_consumer_score_scale(h, scale), "gmm_set_consumer_score_scale");
    const u32 frameOffsets[2] = {0, T}, stateOffsets[2] = {0, u32(finalWeight_.size())};
    const u32 labelOffsets[2] = {0, u32(labelMixture_.size())};
    const u32* arcTarget = arcTarget_.empty() ? 0 : &arcTarget_[0];
    const u32* arcLabel  = arcLabelSlot_.empty() ? 0 : &arcLabelSlot_[0];
    const f32* arcWeight = arcWeight_.empty() ? 0 : &arcWeight_[0];
    const u32* labelMix  = labelMixture_.empty() ? 0 : &labelMixture_[0];
    int        rc;
    if (viterbi_) {
        pathSlot_.assign(std::max<u32>(1, T), 0xFFFFFFFFu);
        float score = Core::Type<Score>::max;
        rc = gmm_score_align_labels_host(h, 1, frameOffsets, x, stateOffsets, &arcOffsets_[0], arcTarget, arcLabel, arcWeight,
                                         &initial_, &finalWeight_[0], labelOffsets, labelMix, &pathSlot_[0], 0, &score);
        score_ = score;
    }
    else {
        const size_t ldW = std::max<size_t>(1, labelMixture_.size());
        weights_.assign(size_t(std::max<u32>(1, T)) * ldW, 0.f);
        double total = Core::Type<Score>::max;
        rc = gmm_score_baum_welch_labels_host(h, 1, frameOffsets, x, stateOffsets, &arcOffsets_[0], arcTarget, arcLabel,
                                              arcWeight, &initial_, &finalWeight_[0], labelOffsets, labelMix, &weights_[0],
                                              ldW, &total);
        // Aligner::alignmentScore() is the f32 forward potential; the device total is the f64 one
        score_ = total >= double(Core::Type<Score>::max) ? Core::Type<Score>::max : Score(total);
    }
    gmm_set_consumer_score_scale(h, 1.f);
    parent->check(rc, viterbi_ ? "gmm_score_align_labels_host" : "gmm_score_baum_welch_labels_host");
    nFrames_ = T;
    ++nCalls_;
}

namespace {
struct ByTimeAndDecreasingWeight {  // Alignment::sortItems(false) on -log weights; equal weights: by label
    bool operator()(const Speech::AlignmentItem& a, const Speech::AlignmentItem& b) const {
        if (a.time != b.time)
            return a.time < b.time;
        if (a.weight != b.weight)
            return a.weight > b.weight;
        return a.emission < b.emission;
    }
};
}  // namespace

void GpuAligner::getAlignment(Speech::Alignment& result) const {
    result.clear();
    if (reachedFinalState()) {
        if (viterbi_) {
            for (u32 t = 0; t < nFrames_; ++t)
                result.push_back(Speech::AlignmentItem(t, labelOfSlot_[pathSlot_[t]], 1.0));
        }
        else {
            const size_t ldW = std::max<size_t>(1, labelMixture_.size());
            for (u32 t = 0; t < nFrames_; ++t)
                for (u32 j = 0; j < labelOfSlot_.size(); ++j)
                    if (weights_[t * ldW + j] > 0.f)  // filterWeightsGT(0.0), Aligner.cc:707
                        result.push_back(Speech::AlignmentItem(t, labelOfSlot_[j], weights_[t * ldW + j]));
            std::sort(result.begin(), result.end(), ByTimeAndDecreasingWeight());
        }
    }
    result.setScore(alignmentScore());
    if (acousticModel_)
        result.setAlphabet(acousticModel_->allophoneStateAlphabet());
}
