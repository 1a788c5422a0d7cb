"""Alignment automata for the device-resident Viterbi consumer (include/gmm_b200.h gmm_align_device).

The reference aligns a segment against an Fsa automaton whose arcs carry an allophone-state label and a transition
weight (Search::Aligner::setModel, /root/reference/src/Search/Aligner.cc:118-122); the label is mapped to an emission
index when the arc is scored (Aligner.cc:201).  Here that mapping is applied once, when the automaton is flattened:
arcs carry the emission index (the column of the score block) directly.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

NOT_FINAL = np.float32(np.inf)


@dataclass
class AlignmentAutomata:
    """Automata of several segments, concatenated.  Arcs of (global) state s: arc_offsets[s]..arc_offsets[s+1];
    targets and initial states are segment-local; final_weight[s] = +inf marks a non-final state."""
    state_offsets: np.ndarray   # [segments + 1] u32
    arc_offsets: np.ndarray     # [states + 1] u32
    arc_target: np.ndarray      # [arcs] u32
    arc_emission: np.ndarray    # [arcs] u32
    arc_weight: np.ndarray      # [arcs] f32
    initial_state: np.ndarray   # [segments] u32
    final_weight: np.ndarray    # [states] f32

    @property
    def segments(self) -> int:
        return self.initial_state.shape[0]

    @classmethod
    def from_arc_lists(cls, automata):
        """automata: one (states, initial, final) triple per segment; states[s] = list of (target, emission, weight)
        in Fsa arc order, final = {state: weight}."""
        so, ao, tg, em, w, ini, fin = [0], [0], [], [], [], [], []
        for states, initial, final in automata:
            for s, arcs in enumerate(states):
                for (t, e, wt) in arcs:
                    tg.append(t)
                    em.append(e)
                    w.append(wt)
                ao.append(len(tg))
                fin.append(final.get(s, NOT_FINAL))
            so.append(so[-1] + len(states))
            ini.append(initial)
        u = lambda a: np.asarray(a, np.uint32)  # noqa: E731
        return cls(u(so), u(ao), u(tg), u(em), np.asarray(w, np.float32), u(ini), np.asarray(fin, np.float32))

    def slice(self, s0: int, s1: int) -> "AlignmentAutomata":
        """Automata of segments [s0, s1) (segment sharding across ranks)."""
        st0, st1 = int(self.state_offsets[s0]), int(self.state_offsets[s1])
        a0, a1 = int(self.arc_offsets[st0]), int(self.arc_offsets[st1])
        return AlignmentAutomata(self.state_offsets[s0:s1 + 1] - np.uint32(st0),
                                 self.arc_offsets[st0:st1 + 1] - np.uint32(a0), self.arc_target[a0:a1],
                                 self.arc_emission[a0:a1], self.arc_weight[a0:a1], self.initial_state[s0:s1],
                                 self.final_weight[st0:st1])


def linear_hmm(emissions, loop=3.0, forward=0.0, skip=30.0, exit_weight=0.0):
    """Left-to-right automaton over a sequence of emission indices with loop / forward / skip transitions, the
    shape RASR's alignment graphs take for one pronunciation (Am::TransitionModel applied to an allophone-state
    chain).  State 0 is the entry (no frame consumed yet); state i+1 = `emissions[i]` was emitted last.
    -> (states, initial, final) for AlignmentAutomata.from_arc_lists."""
    n = len(emissions)
    states = []
    for s in range(n + 1):
        arcs = []
        if s > 0:
            arcs.append((s, int(emissions[s - 1]), float(loop)))
        if s < n:
            arcs.append((s + 1, int(emissions[s]), float(forward)))
        if s + 1 < n and skip is not None:
            arcs.append((s + 2, int(emissions[s + 1]), float(skip)))
        states.append(arcs)
    return states, 0, {n: float(exit_weight)}


def label_slots(automata: AlignmentAutomata, emission_of_label):
    """Arcs that carry LABELS (RASR: allophone states, `arc_emission` holds the label id) for the label entry points
    (gmm_score_align_labels_host / gmm_score_baum_welch_labels_host): per segment the distinct labels in order of first
    use by an arc become slots.  `emission_of_label[label]` = the mixture the acoustic model ties the label to
    (Am::AcousticModel::emissionIndex).
    -> (automata whose arc_emission is the segment-local slot, label_offsets [segments+1], label_ids, label_mixture)."""
    a = automata
    tie = np.asarray(emission_of_label, np.uint32)
    slots = np.empty_like(a.arc_emission)
    offsets, ids = [0], []
    for i in range(a.segments):
        s0, s1 = int(a.state_offsets[i]), int(a.state_offsets[i + 1])
        a0, a1 = int(a.arc_offsets[s0]), int(a.arc_offsets[s1])
        slot_of = {}
        for k in range(a0, a1):
            lab = int(a.arc_emission[k])
            if lab not in slot_of:
                slot_of[lab] = len(slot_of)
                ids.append(lab)
            slots[k] = slot_of[lab]
        offsets.append(len(ids))
    ids = np.asarray(ids, np.uint32)
    out = AlignmentAutomata(a.state_offsets, a.arc_offsets, a.arc_target, slots, a.arc_weight, a.initial_state,
                            a.final_weight)
    return out, np.asarray(offsets, np.uint32), ids, tie[ids] if ids.size else np.zeros(0, np.uint32)
