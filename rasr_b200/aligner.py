"""Host-side mirror of the reference's aligner surface over the C-ABI (tests / bench convenience).

The production host side is the C++ class in rasr/GpuAligner.{hh,cc}; this module states the same interface in Python
so that the parity tests read like calls into the reference's Search::Aligner
(/root/reference/src/Search/Aligner.hh:36-176, driven by /root/reference/src/Speech/AlignmentNode.cc:211, 313-322):

    aligner.setModel(model, acoustic_model)   flatten the automaton; labels -> mixtures (Aligner.cc:225)
    aligner.feed(scorers)                     the context scorers of a segment's frames
    aligner.reachedFinalState(), aligner.alignmentScore()
    aligner.getAlignment()                    items (time, label, weight) as Aligner::getAlignment leaves them

The frames behind the scorers go to the device once; only the emissions the automaton names are scored, the score block
stays in HBM and the search runs there (gmm_score_align_labels_host / gmm_score_baum_welch_labels_host).  Unpruned,
min-weight 0: what the reference computes with acoustic pruning disabled.
"""
from __future__ import annotations

import numpy as np

from .align import AlignmentAutomata, label_slots
from .feature_scorer import ContextScorer, GpuFeatureScorer

FLT_MAX = float(np.finfo(np.float32).max)


class Aligner:
    modeViterbi, modeBaumWelch = "viterbi", "baum-welch"

    def __init__(self, config: dict | None = None):
        config = config or {}
        self._mode = config.get("mode", self.modeViterbi)             # Aligner.cc:498-499
        if self._mode not in (self.modeViterbi, self.modeBaumWelch):
            raise ValueError(f"unknown aligner mode '{self._mode}'")
        if float(config.get("min-weight", 0.0)) > 0.0:                 # Aligner.cc:501-504, 564-571
            raise ValueError("the GPU aligner keeps every alignment item: min-weight must be 0")
        self._model = None
        self.restart()

    def selectMode(self, mode: str) -> None:
        if mode not in (self.modeViterbi, self.modeBaumWelch):
            raise ValueError(f"unknown aligner mode '{mode}'")
        self._mode = mode

    def getMode(self) -> str:
        return self._mode

    def setModel(self, model, acoustic_model) -> None:
        """model: (states, initial, final) with states[s] = [(target, label, weight), ...] in Fsa arc order and
        final = {state: weight}; acoustic_model: label -> mixture (array or callable, Am::AcousticModel::emissionIndex)."""
        states, initial, final = model
        assert 0 <= initial < max(1, len(states)), "model needs an initial state"   # require(), Aligner.cc:575-576
        au = AlignmentAutomata.from_arc_lists([model])
        n_labels = int(au.arc_emission.max()) + 1 if au.arc_emission.size else 0
        tie = (np.asarray([acoustic_model(l) for l in range(n_labels)], np.uint32) if callable(acoustic_model)
               else np.asarray(acoustic_model, np.uint32))
        self._model = label_slots(au, tie)     # (automata with slots, label_offsets, label_ids, label_mixture)
        self._initial_final = final.get(initial)
        self.restart()

    def restart(self) -> None:
        self._score, self._items = FLT_MAX, []

    def feed(self, scorers, scale: float = 1.0) -> None:
        """scorers: the ContextScorer of every frame, from ONE GpuFeatureScorer on the float engine; `scale` is what
        Mm::FeatureScorerScaling would multiply every score by (ScaledFeatureScorer.hh:62-64)."""
        assert self._model is not None, "setModel first"
        self.restart()
        T = len(scorers)
        if T == 0:   # the start hypothesis is the only one (Aligner.cc:376-396)
            if self._initial_final is not None:
                self._score = float(np.float32(self._initial_final))
            return
        parent = scorers[0]._parent
        for c in scorers:
            if not isinstance(c, ContextScorer) or c._parent is not parent:
                raise ValueError("the GPU aligner needs all frames of a segment from one GPU feature scorer")
            if c._frame is None:
                raise ValueError("the scorer released its frame (scored before the aligner saw it): create the feature "
                                 "scorer with keep_frames=True")
        if parent._type in GpuFeatureScorer.QUANTIZED_TYPES:
            raise ValueError("the GPU aligner runs on the float engine")
        x = np.stack([c._frame for c in scorers]).astype(np.float32)
        slots, lo, ids, mix = self._model
        eng = parent.engine
        eng.set_consumer_score_scale(scale)
        try:
            if self._mode == self.modeViterbi:
                lab, _, sc = eng.score_align_labels_host([0, T], x, slots, lo, mix, want_states=False)
                self._score = float(sc[0])
                if sc[0] != np.float32(FLT_MAX):
                    self._items = [(t, int(ids[lab[t]]), 1.0) for t in range(T)]
            else:
                w, total = eng.score_baum_welch_labels_host([0, T], x, slots, lo, mix)
                self._score = FLT_MAX if total[0] >= FLT_MAX else float(np.float32(total[0]))
                if self._score != FLT_MAX:
                    for t in range(T):
                        row = [(t, int(ids[j]), float(w[t, j])) for j in range(len(ids)) if w[t, j] > 0]   # Aligner.cc:707
                        self._items += sorted(row, key=lambda it: (-it[2], it[1]))                        # sortItems(false)
        finally:
            eng.set_consumer_score_scale(1.0)

    def alignmentScore(self) -> float:
        return self._score

    def reachedFinalState(self) -> bool:
        return self._score != FLT_MAX

    def getAlignment(self):
        """[(time, label, weight)]: weight 1 per frame in Viterbi mode; in Baum-Welch mode every positive normalised
        state posterior, by time and decreasing weight."""
        return list(self._items)
