"""Generates tests/golden/aln_*.npz: the UNMODIFIED reference's Search::Aligner (src/Search/Aligner.cc, with its Fsa
library and Speech::Alignment) run by oracle/_ref/ref_aligner (oracle/ref_build/ref_aligner.cc) in both modes, unpruned,
over synthetic score blocks and automata.  Run in the build container only:  python tests/golden/make_align_golden.py

Each fixture holds the inputs -- frame_offsets, scores [frames x emissions] f32, the automata arrays of
rasr_b200.align.AlignmentAutomata -- and what the reference answered per segment:
  vit_emission [frames] u32   the emission of the alignment item at each frame (0xFFFFFFFF: no final state reached)
  vit_score [segments] f32    Aligner::alignmentScore() in mode viterbi
  bw_item_offsets, bw_time (segment-local), bw_emission, bw_weight f64   the items of Speech::Alignment in mode baum-welch
  bw_score [segments] f32     Aligner::alignmentScore() in mode baum-welch (f32 log-semiring forward potential)
  bw_total_inv [segments] f32 totalInv of Fsa::posterior64: MINUS the f64 forward-backward total, rounded once
aln_tied also holds emission_of_label: its arcs carry labels (arc_emission = label id, vit_emission / bw_emission name
labels) which the acoustic model ties to the columns of `scores`."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from rasr_b200.align import AlignmentAutomata, linear_hmm  # noqa: E402


def random_graph(rng, S, M, integer_weights):
    """General automaton: random out-degree, random targets (cycles, dead ends), several final states."""
    states = []
    for _ in range(S):
        arcs = []
        for _ in range(int(rng.integers(0, 5))):
            w = float(rng.integers(0, 3)) if integer_weights else float(rng.uniform(0, 5))
            arcs.append((int(rng.integers(0, S)), int(rng.integers(0, M)), w))
        states.append(arcs)
    final = {int(s): (float(rng.integers(0, 2)) if integer_weights else float(rng.uniform(0, 3)))
             for s in rng.choice(S, size=max(1, S // 4), replace=False)}
    return states, int(rng.integers(0, S)), final


def case_linear():
    rng = np.random.default_rng(101)
    M, lens = 48, [1, 2, 7, 30, 61, 18]
    segs = [linear_hmm(rng.integers(0, M, max(1, n // int(rng.integers(2, 5)))), loop=float(rng.uniform(0, 4)),
                       forward=float(rng.uniform(0, 2)), skip=float(rng.uniform(2, 30))) for n in lens]
    return lens, rng.uniform(20, 120, (sum(lens), M)).astype(np.float32), segs


def case_graphs():
    rng = np.random.default_rng(102)
    M, lens = 32, [5, 12, 1, 25, 40, 9, 16, 3]
    segs = [random_graph(rng, int(rng.integers(2, 30)), M, False) for _ in lens]
    return lens, rng.uniform(0, 60, (sum(lens), M)).astype(np.float32), segs


def case_ties():
    """Integer scores and weights: exact f32 ties everywhere, resolved by the reference's hypothesis order."""
    rng = np.random.default_rng(103)
    M, lens = 6, [4, 9, 15, 2, 22, 7, 11, 30, 5, 13]
    segs = [random_graph(rng, int(rng.integers(2, 14)), M, True) for _ in lens]
    segs[0] = ([[(1, 3, 0.0), (2, 4, 0.0)], [(3, 5, 0.0)], [(3, 0, 0.0)], [(3, 1, 0.0)]], 0, {3: 0.0})
    lens[0] = 2
    segs[3] = ([[(2, 1, 0.0), (1, 0, 0.0)], [], []], 0, {1: 0.0, 2: 0.0})
    lens[3] = 1
    return lens, rng.integers(0, 3, (sum(lens), M)).astype(np.float32), segs


def case_edges():
    """No final state reachable; a single frame; one state looping on itself; a wide fan; segments without frames
    (the start hypothesis is then the last frame's only hypothesis: final or not)."""
    rng = np.random.default_rng(104)
    M = 8
    fan = ([[(1 + k, k, float(k)) for k in range(7)]] + [[(8, (k + 1) % M, 0.5)] for k in range(7)] + [[]], 0, {8: 1.0})
    segs = [linear_hmm([1, 2, 3, 4, 5]), linear_hmm([7]), ([[(0, 2, 0.25)]], 0, {0: 2.5}), fan,
            ([[(1, 0, 0.0)], []], 0, {1: 0.0}), ([[(0, 0, 0.0)]], 0, {0: 2.5}), ([[(1, 0, 0.0)], []], 0, {1: 0.0})]
    lens = [2, 1, 6, 2, 3, 0, 0]
    return lens, rng.uniform(0, 9, (sum(lens), M)).astype(np.float32), segs


def case_long():
    rng = np.random.default_rng(105)
    M, lens = 64, [400]
    segs = [linear_hmm(rng.integers(0, M, 150), loop=3.0, forward=0.0, skip=30.0)]
    return lens, rng.uniform(20, 120, (400, M)).astype(np.float32), segs


def case_tied():
    """State tying: arcs carry LABELS (allophone states), the acoustic model maps several labels to one emission, and
    the reference's items name the label (Aligner.cc:185 / :205 store a->input_).  Parallel arcs between the same two
    states whose labels share an emission, integer weights in some segments (ties between such arcs)."""
    rng = np.random.default_rng(106)
    M, n_labels, lens = 12, 40, [6, 14, 3, 25, 9, 18, 1, 11]
    tie = rng.integers(0, M, n_labels).astype(np.uint32)
    segs = []
    for k, _ in enumerate(lens):
        states, ini, fin = random_graph(rng, int(rng.integers(2, 16)), n_labels, integer_weights=bool(k & 1))
        for arcs in states:                       # duplicate some arcs under another label of the same emission
            for (tg, lab, w) in list(arcs):
                same = np.flatnonzero(tie == tie[lab])
                if same.size > 1 and rng.random() < 0.3:
                    arcs.append((tg, int(rng.choice(same)), w if rng.random() < 0.5 else w + 1.0))
        segs.append((states, ini, fin))
    scores = np.where(rng.random((sum(lens), M)) < 0.5, rng.integers(0, 4, (sum(lens), M)),
                      rng.uniform(0, 4, (sum(lens), M))).astype(np.float32)
    return lens, scores, segs, tie


CASES = {"aln_tied": case_tied, "aln_linear": case_linear, "aln_graphs": case_graphs, "aln_ties": case_ties, "aln_edges": case_edges,
         "aln_long": case_long}


def main():
    assert O.build_ref() and O.ref_aligner_available(), "needs /root/reference (build container)"
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, make in CASES.items():
        lens, scores, segs, *tie = make()
        tie = tie[0] if tie else None
        au = AlignmentAutomata.from_arc_lists(segs)
        fo = np.cumsum([0] + list(lens)).astype(np.uint32)
        vit_em = np.full(int(fo[-1]), 0xFFFFFFFF, np.uint32)
        vit_sc, bw_sc, bw_tot = (np.empty(len(lens), np.float32) for _ in range(3))
        io, it, ie, iw = [0], [], [], []
        for i in range(len(lens)):
            rows = scores[fo[i]:fo[i + 1]]
            v = O.run_ref_aligner(rows, au, i, mode="viterbi", emission_of_label=tie)
            vit_sc[i] = v["score"]
            if v["time"].size:
                assert np.array_equal(np.sort(v["time"]), np.arange(lens[i])) and np.all(v["weight"] == 1.0)
                vit_em[int(fo[i]) + v["time"]] = v["emission"]
            b = O.run_ref_aligner(rows, au, i, mode="baum-welch", emission_of_label=tie)
            bw_sc[i], bw_tot[i] = b["score"], b["total"]
            it.append(b["time"]), ie.append(b["emission"]), iw.append(b["weight"])
            io.append(io[-1] + b["time"].size)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), frame_offsets=fo, scores=scores,
                            state_offsets=au.state_offsets, arc_offsets=au.arc_offsets, arc_target=au.arc_target,
                            arc_emission=au.arc_emission, arc_weight=au.arc_weight, initial_state=au.initial_state,
                            final_weight=au.final_weight, vit_emission=vit_em, vit_score=vit_sc,
                            bw_item_offsets=np.asarray(io, np.uint32), bw_time=np.concatenate(it).astype(np.uint32),
                            bw_emission=np.concatenate(ie).astype(np.uint32), bw_weight=np.concatenate(iw),
                            bw_score=bw_sc, bw_total_inv=bw_tot,
                            **({"emission_of_label": tie} if tie is not None else {}))
        print(name, "ok: segments", len(lens), "frames", int(fo[-1]), "baum-welch items", io[-1], "viterbi scores", vit_sc[:4])


if __name__ == "__main__":
    main()
