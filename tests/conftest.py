import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_names():
    """Model fixtures (scorer goldens); lda_*.npz (front end) and est_*.npz (estimators, model files) are fixtures
    with their own tests."""
    names = (os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
    return sorted(n for n in names if not n.startswith(("lda_", "est_", "batchint_", "aln_")))


def batchint_golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "batchint_*.npz")))


def align_golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "aln_*.npz")))


def load_align_golden(name):
    """-> (frame_offsets, scores, AlignmentAutomata, fixture): what the reference's Search::Aligner was given and what
    it answered (tests/golden/make_align_golden.py)."""
    from rasr_b200.align import AlignmentAutomata
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    au = AlignmentAutomata(z["state_offsets"], z["arc_offsets"], z["arc_target"], z["arc_emission"], z["arc_weight"],
                           z["initial_state"], z["final_weight"])
    return z["frame_offsets"], z["scores"], au, z


def estimator_golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "est_*.npz")))


ESTIMATOR_CASES = {"vit": ("max", True, False), "vitw": ("max", True, True), "vitsum": ("sum", True, True),
                   "bw": ("sum", False, True)}   # case -> (scorer, viterbi, weighted); tests/golden/make_estimator_golden.py


def load_golden(name):
    from rasr_b200.gms import MixtureSetData
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    ms = MixtureSetData(int(z["D"]), z["mix_offsets"], z["mix_density"], z["mix_logw"], z["dens_mean"], z["dens_cov"],
                        z["means"], z["variances"])
    return ms, z


def ref_quant_tables(ms, z):
    """The quantized tables exactly as the compiled reference's init() produced them (from a fixture)."""
    return dict(offsets=ms.mix_offsets, qmean=z["simd_qmean"], cov_idx=z["simd_qcov"], qconst=z["simd_qconst"],
                isr_scaled=z["simd_isrs"], scaling_squared=np.float32(z["simd_scaling_squared"]))


def oracle_float_tables(ms, **kw):
    """Float tables with the reference build's rsqrt-based inverse std-dev (oracle restatement)."""
    from oracle import oracle as O
    from rasr_b200 import tables
    isr, logn, w2 = O.float_tables(ms, **kw)
    mean, cov_idx = tables.flatten(ms)
    return dict(offsets=ms.mix_offsets, mean=mean, cov_idx=cov_idx, isr=isr, log_norm=logn, w2=w2)


def oracle_batchint_tables(ms, exact_rsqrt=False):
    """Tables of the reference's batch-diagonal-maximum-int scorer (oracle restatement), in upload_quant's layout."""
    from oracle import oracle as O
    t = O.batchint_prepare(ms, exact_rsqrt=exact_rsqrt)
    return dict(offsets=ms.mix_offsets, qmean=t["qmean"], cov_idx=t["qcov"], qconst=t["qconst"], isr_scaled=t["isrs"],
                scaling_squared=np.float32(t["scale2"] * np.float32(0.5))), t


def oracle_quant_tables(ms, exact_rsqrt=False):
    from oracle import oracle as O
    t = O.simd_prepare(ms, exact_rsqrt=exact_rsqrt)
    return dict(offsets=ms.mix_offsets, qmean=t["qmean"], cov_idx=t["qcov"], qconst=t["qconst"],
                isr_scaled=t["isrs"], scaling_squared=t["scaling_squared"]), t


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O
