"""Build container only (needs oracle/_ref/ref_aligner = the reference's own Search::Aligner): how far is the reference's
DEFAULT search -- acoustic pruning at 50, doubled until the alignment score stops changing (Aligner.cc:462-481,584-611)
-- from the unpruned search MmGpu::GpuAligner and the device consumers implement?  Left-to-right acceptors over random
labels, two kinds of score blocks: uniform random scores, and the scores of a synthetic mixture set on its own features.
    python profiles/ref_pruning_vs_unpruned.py > profiles/ref_pruning_vs_unpruned_r02.txt"""
import os
import subprocess
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402
from rasr_b200 import synth  # noqa: E402
from rasr_b200.align import AlignmentAutomata, linear_hmm  # noqa: E402


def run(scores, au, mode, pruned):
    with tempfile.TemporaryDirectory() as tmp:
        ip, op = os.path.join(tmp, "in.bin"), os.path.join(tmp, "out")
        s1 = int(au.state_offsets[1])
        a1 = int(au.arc_offsets[s1])
        with open(ip, "wb") as f:
            np.asarray([scores.shape[0], scores.shape[1], s1, a1, int(au.initial_state[0])], np.uint32).tofile(f)
            for a in (au.arc_offsets[:s1 + 1], au.arc_target[:a1], au.arc_emission[:a1], au.arc_weight[:a1],
                      au.final_weight[:s1], scores):
                np.ascontiguousarray(a).tofile(f)
        cmd = [O.REF_ALIGNER_BIN, f"in={ip}", f"out={op}", f"mode={mode}"]
        if not pruned:
            cmd += ["--*.aligner.min-acoustic-pruning=1e30", "--*.aligner.max-acoustic-pruning=1e30"]
        r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert r.returncode == 0, r.stdout[-500:]
        raw = open(op + ".items.bin", "rb").read()
        n = int(np.frombuffer(raw[:4], np.uint32)[0])
        items = np.frombuffer(raw[4:], np.dtype([("t", "<u4"), ("e", "<u4"), ("w", "<f8")]), count=n)
        return items, np.fromfile(op + ".score.f32", np.float32)[0]


def experiment(name, make_scores, trials=30):
    rng = np.random.default_rng(5)
    same = worse = 0
    dws = []
    for _ in range(trials):
        T = int(rng.integers(50, 400))
        scores = make_scores(rng, T)
        au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, scores.shape[1], max(1, T // 3)), loop=3.0,
                                                          forward=0.0, skip=30.0)])
        a, sa = run(scores, au, "viterbi", True)
        b, sb = run(scores, au, "viterbi", False)
        same += int(np.array_equal(a["e"], b["e"]) and sa.view(np.uint32) == sb.view(np.uint32))
        worse += int(sa > sb)
        a, _ = run(scores, au, "baum-welch", True)
        b, _ = run(scores, au, "baum-welch", False)
        da = {(int(t), int(e)): w for t, e, w in a}
        db = {(int(t), int(e)): w for t, e, w in b}
        dws.append(max(abs(da.get(k, 0) - db.get(k, 0)) for k in set(da) | set(db)))
    print(f"{name}: viterbi default-pruned == unpruned (items and score bits) in {same} of {trials} segments "
          f"(pruned score worse in {worse}); baum-welch max |weight difference| per segment: median "
          f"{np.median(dws):.2g}, max {np.max(dws):.2g}")


def main():
    assert O.ref_aligner_available()
    experiment("uniform random scores in [60, 160)", lambda rng, T: rng.uniform(60, 160, (T, 200)).astype(np.float32))
    ms = synth.make_model(M=200, K=8, D=24, seed=3)

    def gmm_scores(rng, T):
        x = synth.make_features(ms, T, seed=int(rng.integers(1 << 30)))
        return O.score_max(ms, x)[0]
    experiment("scores of a 200 x 8 x 24 mixture set on synthetic features", gmm_scores)


if __name__ == "__main__":
    main()
