// TEST INFRASTRUCTURE (oracle/): an Am::AcousticModel whose state tying is a table, for the harnesses around the
// reference's Search::Aligner (ref_aligner.cc, rasr/check.cc).  The aligner asks its acoustic model for
// emissionIndex(label) (Aligner.cc:225) and for the alphabet it attaches to the result (Aligner.cc:718) -- nothing else.
#pragma once
#include <cstdlib>
#include <vector>

#include <Am/AcousticModel.hh>

class TableAcousticModel : public Am::AcousticModel {
    u32              n_;
    std::vector<u32> emissionOfLabel_;  // empty: labels are emission indices

public:
    TableAcousticModel(const Core::Configuration& c, u32 n, const std::vector<u32>& emissionOfLabel)
            : Core::Component(c), Am::AcousticModel(c), n_(n), emissionOfLabel_(emissionOfLabel) {}
    virtual void load(Mode) {}
    virtual void getDependencies(Core::DependencySet&) const {}
    virtual Core::Ref<Am::TransducerBuilder> createTransducerBuilder() const {
        return Core::Ref<Am::TransducerBuilder>();
    }
    virtual Core::Ref<const Bliss::PhonemeInventory> phonemeInventory() const {
        return Core::Ref<const Bliss::PhonemeInventory>();
    }
    virtual Core::Ref<const Am::AllophoneAlphabet> allophoneAlphabet() const {
        return Core::Ref<const Am::AllophoneAlphabet>();
    }
    virtual Core::Ref<const Am::AllophoneStateAlphabet> allophoneStateAlphabet() const {
        return Core::Ref<const Am::AllophoneStateAlphabet>();
    }
    virtual Core::Ref<const Am::Phonology> phonology() const {
        return Core::Ref<const Am::Phonology>();
    }
    virtual Bliss::Phoneme::Id silence() const {
        return 0;
    }
    virtual Am::AllophoneStateIndex silenceAllophoneStateIndex() const {
        return 0;
    }
    virtual Core::Ref<Mm::AbstractMixtureSet> mixtureSet() {
        return Core::Ref<Mm::AbstractMixtureSet>();
    }
    virtual Core::Ref<const Mm::ScaledFeatureScorer> featureScorer() {
        return Core::Ref<const Mm::ScaledFeatureScorer>();
    }
    virtual Core::Ref<Mm::ScaledFeatureScorer> mutableFeatureScorer() {
        return Core::Ref<Mm::ScaledFeatureScorer>();
    }
    virtual bool setFeatureScorer(Core::Ref<Mm::ScaledFeatureScorer>) {
        return false;
    }
    virtual EmissionIndex emissionIndex(Am::AllophoneState) const {
        abort();
    }
    virtual EmissionIndex emissionIndex(Am::AllophoneStateIndex i) const {
        return emissionOfLabel_.empty() ? i : emissionOfLabel_[i];  // state tying: several labels, one emission
    }
    virtual EmissionIndex nEmissions() const {
        return n_;
    }
    virtual StateTransitionIndex nStateTransitions() const {
        return 0;
    }
    virtual const Am::StateTransitionModel* stateTransition(StateTransitionIndex) const {
        return 0;
    }
    virtual StateTransitionIndex stateTransitionIndex(Am::AllophoneState, s8) const {
        return 0;
    }
    virtual StateTransitionIndex stateTransitionIndex(Am::AllophoneStateIndex, s8) const {
        return 0;
    }
    virtual Core::Ref<const Am::ClassicHmmTopologySet> hmmTopologySet() const {
        return Core::Ref<const Am::ClassicHmmTopologySet>();
    }
    virtual const Am::ClassicHmmTopology* hmmTopology(Bliss::Phoneme::Id) const {
        return 0;
    }
    virtual bool isAcrossWordModelEnabled() const {
        return false;
    }
};
