"""GPU: the RASR-side plugin classes (rasr/GpuFeatureScorer.cc) against the reference's own scorer classes, both
created through Mm::FeatureScorerFactory inside one RASR process and driven by the buffered caller loop of
Speech::AlignmentNode.  The binary (oracle/_ref/rasr_gpu_check) is built by rasr/Makefile in the build container
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
    assert r.stdout.count("PASS") == 4
