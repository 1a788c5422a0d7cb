"""GPU: the RASR-side plugin classes (rasr/GpuFeatureScorer.cc, rasr/GpuAligner.cc) against the reference's own scorer
classes and its own Search::Aligner: the scorers are
created by the reference's own Mm::Module_ (feature-scorer-type -> factory -> scaling wrapper; rasr/Module.cc registers
the GPU scorers in it) inside one RASR process and driven by the buffered caller loops of Speech::AlignmentNode and
Speech::RecognizerDelayHandler.  The binary (oracle/_ref/rasr_gpu_check) is built by rasr/Makefile in the build container
from the unmodified reference objects; it travels to the GPU box with the snapshot."""
import os
import subprocess

import pytest

from conftest import ROOT
from rasr_b200 import synth

pytestmark = pytest.mark.gpu
CHECK = os.path.join(ROOT, "oracle", "_ref", "rasr_gpu_check")


@pytest.mark.skipif(not os.path.exists(CHECK), reason="oracle/_ref/rasr_gpu_check not built (needs /root/reference)")
@pytest.mark.parametrize("kw,T,buffer_size", [
    (dict(M=301, K=16, D=45), 300, 0),
    (dict(M=301, K=16, D=45), 300, 32),
    (dict(M=64, K=64, D=45, seed=2), 150, 0),
    (dict(M=90, K=7, D=33, covariance="density", ragged=True, shared=True, zero_weight_frac=0.2, seed=3), 100, 8),
])
def test_plugin_matches_reference_scorers_in_process(tmp_path, kw, T, buffer_size):
    ms = synth.make_model(**kw)
    x = synth.make_features(ms, T)
    mp, fp = str(tmp_path / "m.gms"), str(tmp_path / "x.f32")
    ms.save(mp)
    x.tofile(fp)
    r = subprocess.run([CHECK, f"model={mp}", f"features={fp}", f"buffer-size={buffer_size}"], cwd=str(tmp_path),
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    print(r.stdout)
    assert r.returncode == 0, r.stdout[-3000:]
    # 7 scorer pairs (the three batch scorers need a pooled covariance), delayed-recognizer loop, time-indexed scorers,
    # density posteriors, model accessors, with-density-assignment, and MmGpu::GpuAligner against the reference's own
    # Search::Aligner in both modes on both float kernels (4 lines)
    assert r.stdout.count("PASS") == (13 if kw.get("covariance") else 16) and "FAIL" not in r.stdout


@pytest.mark.skipif(not os.path.exists(CHECK), reason="oracle/_ref/rasr_gpu_check not built (needs /root/reference)")
def test_plugin_path_is_within_20_percent_of_the_call_it_wraps(tmp_path):
    """BASELINE.json configs[0]: 4501 x 16 x 45, one 1000-frame segment; segments of 4000 frames for the timing.  The
    plugin path (addFeature per frame -> one batch -> flush + one read per frame) against gmm_score_host on the same
    engine, best of 12 repetitions each, in the same process."""
    ms = synth.make_model(M=4501, K=16, D=45)
    x = synth.make_features(ms, 64)
    mp, fp = str(tmp_path / "m.gms"), str(tmp_path / "x.f32")
    ms.save(mp)
    x.tofile(fp)
    r = subprocess.run([CHECK, f"model={mp}", f"features={fp}", "buffer-size=0", "timing-frames=4000"], cwd=str(tmp_path),
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    print(r.stdout)
    assert r.returncode == 0, r.stdout[-3000:]
    import re
    timing = re.findall(r"TIMING frames=\d+ mixtures=\d+ plugin_frames_per_s=\d+ gmm_score_host_frames_per_s=\d+ ratio=([0-9.]+)",
                        r.stdout)  # (RASR's own log lines share stdout and may end without a newline)
    assert len(timing) == 1, r.stdout[-2000:]
    assert float(timing[0]) >= 0.8, timing
    assert r.stdout.count("PASS") == 18 and "FAIL" not in r.stdout
