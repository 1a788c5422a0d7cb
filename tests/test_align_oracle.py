"""CPU: the oracle's Viterbi restatement (oracle/gmm_oracle.c orc_viterbi_align) against an independent dense DP,
and the hypothesis-order tie rules of /root/reference/src/Search/Aligner.cc:132-158, 398-411 on hand-made cases."""
import numpy as np
import pytest

from rasr_b200.align import AlignmentAutomata, linear_hmm


def dense_viterbi_score(scores, states, initial, final):
    """Best final score by a dense float32 DP over (frame, state); no path, no tie rules."""
    S, T = len(states), scores.shape[0]
    cur = np.full(S, np.inf, np.float32)
    cur[initial] = 0
    for t in range(T):
        nxt = np.full(S, np.inf, np.float32)
        for s, arcs in enumerate(states):
            if not np.isfinite(cur[s]):
                continue
            for (tg, em, w) in arcs:
                c = np.float32(np.float32(cur[s] + np.float32(w)) + scores[t, em])
                if c < nxt[tg]:
                    nxt[tg] = c
        cur = nxt
    best = np.float32(np.finfo(np.float32).max)
    for s, w in final.items():
        if np.isfinite(cur[s]):
            best = min(best, np.float32(np.float32(w) + cur[s]))
    return best


def path_cost(scores, states, initial, final, path_state, path_emission):
    """Re-walk a reported path: every step must be an arc of the automaton; returns its f32 cost."""
    s, c = initial, np.float32(0)
    for t in range(scores.shape[0]):
        arcs = [(tg, em, w) for (tg, em, w) in states[s] if tg == path_state[t] and em == path_emission[t]]
        assert arcs, f"frame {t}: no arc {s} -> {path_state[t]} emitting {path_emission[t]}"
        w = min(a[2] for a in arcs)
        c = np.float32(np.float32(c + np.float32(w)) + scores[t, path_emission[t]])
        s = int(path_state[t])
    assert s in final
    return np.float32(np.float32(final[s]) + c)


@pytest.mark.parametrize("seed", range(6))
def test_oracle_viterbi_matches_dense_dp(oracle, seed):
    rng = np.random.default_rng(seed)
    M = 40
    segs, fo = [], [0]
    for _ in range(5):
        n = int(rng.integers(1, 25))
        T = int(rng.integers(max(1, (n + 1) // 2), 3 * n + 4))
        segs.append(linear_hmm(rng.integers(0, M, n), loop=rng.uniform(0, 4), forward=rng.uniform(0, 2),
                               skip=rng.uniform(2, 8)))
        fo.append(fo[-1] + T)
    scores = rng.uniform(0, 50, size=(fo[-1], M)).astype(np.float32)
    au = AlignmentAutomata.from_arc_lists(segs)
    em, st, sc = oracle.viterbi_align(fo, scores, au)
    for i, (states, ini, fin) in enumerate(segs):
        rows = scores[fo[i]:fo[i + 1]]
        want = dense_viterbi_score(rows, states, ini, fin)
        assert sc[i] == want
        if want < np.finfo(np.float32).max:
            assert path_cost(rows, states, ini, fin, st[fo[i]:fo[i + 1]], em[fo[i]:fo[i + 1]]) == want
        else:
            assert np.all(em[fo[i]:fo[i + 1]] == 0xFFFFFFFF)


def test_oracle_viterbi_unreachable_and_empty(oracle):
    # 5 emissions, 2 frames, skips allowed: at most 4 states advanced -> final state 5 not reachable
    a = AlignmentAutomata.from_arc_lists([linear_hmm([1, 2, 3, 4, 5]), linear_hmm([7])])
    scores = np.ones((3, 8), np.float32)
    em, st, sc = oracle.viterbi_align([0, 2, 3], scores, a)
    assert sc[0] == np.finfo(np.float32).max and np.all(em[:2] == 0xFFFFFFFF) and np.all(st[:2] == 0xFFFFFFFF)
    assert sc[1] == 1.0 and em[2] == 7 and st[2] == 1
    # a segment without frames: the start hypothesis is the last frame's only hypothesis
    b = AlignmentAutomata.from_arc_lists([([[(0, 0, 0.0)]], 0, {0: 2.5}), ([[(1, 0, 0.0)], []], 0, {1: 0.0})])
    em, st, sc = oracle.viterbi_align([0, 0, 0], np.zeros((0, 1), np.float32), b)
    assert sc[0] == 2.5 and sc[1] == np.finfo(np.float32).max and em.size == 0


def test_oracle_viterbi_tie_rules(oracle):
    """Equal candidates: the LAST one in hypothesis/arc order wins a state (`sh.score >= score`), the FIRST
    final hypothesis in activation order wins the segment (strict <)."""
    # state 0 -> {1 emitting 3, 2 emitting 4}; 1 -> 3 emitting 5; 2 -> 3 emitting 6; all scores equal.
    states = [[(1, 3, 0.0), (2, 4, 0.0)], [(3, 5, 0.0)], [(3, 6, 0.0)], []]
    a = AlignmentAutomata.from_arc_lists([(states, 0, {3: 0.0})])
    em, st, sc = oracle.viterbi_align([0, 2], np.ones((2, 8), np.float32), a)
    assert list(em) == [4, 6] and list(st) == [2, 3] and sc[0] == 2.0   # later arrival (via state 2) replaces
    # activation order, not state numbering, decides between final states: arcs list target 2 before target 1
    states = [[(2, 1, 0.0), (1, 0, 0.0)], [], []]
    a = AlignmentAutomata.from_arc_lists([(states, 0, {1: 0.0, 2: 0.0})])
    em, st, sc = oracle.viterbi_align([0, 1], np.ones((1, 2), np.float32), a)
    assert list(st) == [2] and list(em) == [1]


def test_automata_slice_is_segment_sharding(oracle):
    """Segments are independent: aligning a slice of the automata over its rows gives the same paths."""
    rng = np.random.default_rng(9)
    segs, fo = [], [0]
    for _ in range(9):
        n = int(rng.integers(1, 12))
        segs.append(linear_hmm(rng.integers(0, 20, n)))
        fo.append(fo[-1] + int(rng.integers(n, 2 * n + 3)))
    fo = np.array(fo, np.uint32)
    scores = rng.uniform(0, 9, size=(fo[-1], 20)).astype(np.float32)
    au = AlignmentAutomata.from_arc_lists(segs)
    em, st, sc = oracle.viterbi_align(fo, scores, au)
    for (s0, s1) in ((0, 4), (4, 9), (3, 3)):
        sub = au.slice(s0, s1)
        e2, t2, c2 = oracle.viterbi_align(fo[s0:s1 + 1] - fo[s0], scores[fo[s0]:fo[s1]], sub)
        assert np.array_equal(e2, em[fo[s0]:fo[s1]]) and np.array_equal(t2, st[fo[s0]:fo[s1]])
        assert np.array_equal(c2, sc[s0:s1])


def _all_paths_posteriors(T, scores, states, initial, final):
    """Brute force: enumerate every path of T arcs from `initial` to a final state; -> (posterior[t][emission], total)."""
    import itertools
    import math
    paths = []

    def walk(s, t, cost, ems):
        if t == T:
            if s in final:
                paths.append((cost + float(np.float32(final[s])), tuple(ems)))
            return
        for (tg, e, w) in states[s]:
            walk(tg, t + 1, cost + float(np.float32(np.float32(w) + scores[t, e])), ems + [e])

    walk(initial, 0, 0.0, [])
    if not paths:
        return None, None
    z = sum(math.exp(-c) for c, _ in paths)
    post = [dict() for _ in range(T)]
    for c, ems in paths:
        for t, e in enumerate(ems):
            post[t][e] = post[t].get(e, 0.0) + math.exp(-c) / z
    return post, -math.log(z)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_oracle_baum_welch_matches_path_enumeration(oracle, seed):
    """The forward-backward restatement against a sum over all paths on automata small enough to enumerate."""
    rng = np.random.default_rng(seed)
    M, T, S = 9, 6, 5
    states = []
    for s in range(S):
        states.append([(int(rng.integers(0, S)), int(rng.integers(0, M)), float(rng.uniform(0, 3)))
                       for _ in range(int(rng.integers(1, 4)))])
    final = {int(s): float(rng.uniform(0, 2)) for s in rng.choice(S, size=2, replace=False)}
    scores = rng.uniform(0, 6, (T, M)).astype(np.float32)
    au = AlignmentAutomata.from_arc_lists([(states, 0, final)])
    w, lists, total = oracle.baum_welch_align(np.array([0, T], np.uint32), scores, au)
    post, tot = _all_paths_posteriors(T, scores, states, 0, final)
    if post is None:
        assert total[0] == np.finfo(np.float32).max and not w[0].any()
        return
    assert abs(total[0] - tot) < 1e-9
    for t in range(T):
        np.testing.assert_allclose(w[0][t].sum(), 1.0, rtol=1e-6)
        for j, e in enumerate(lists[0]):
            assert abs(w[0][t, j] - post[t].get(int(e), 0.0)) < 1e-6


def test_oracle_baum_welch_unreachable(oracle):
    states = [[(1, 0, 0.0)], []]                       # one arc, final state 1: exactly one frame fits
    au = AlignmentAutomata.from_arc_lists([(states, 0, {1: 0.0})])
    scores = np.ones((3, 2), np.float32)
    w, lists, total = oracle.baum_welch_align(np.array([0, 3], np.uint32), scores, au)
    assert total[0] == np.finfo(np.float32).max and not w[0].any()
    w, lists, total = oracle.baum_welch_align(np.array([0, 1], np.uint32), scores[:1], au)
    assert abs(total[0] - 1.0) < 1e-12 and w[0][0, 0] == 1.0
