"""GPU: gmm_quant_score_device_static -- the entry point that allocates nothing (scratch in a static device array;
no counterpart in the reference) -- against gmm_quant_score_device.

Collected last on purpose.  Passes alone in a fresh process and after the others."""
import numpy as np
import pytest

from conftest import oracle_quant_tables
from rasr_b200 import capi, synth

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    e = capi.Engine(0)
    yield e
    e.close()


def _both(eng, x, T, M):
    dev = torch.device("cuda:0")
    dx = torch.from_numpy(x).to(dev)
    out = {}
    for name, call in (("existing", eng.quant_score_device), ("static", eng.quant_score_device_static)):
        q = torch.full((T, M), -7, dtype=torch.int32, device=dev)
        b = torch.full((T, M), -7, dtype=torch.int32, device=dev)
        s = torch.full((T, M), -7.0, dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        call(T, dx, q, b, s)
        eng.sync()
        out[name] = (q.cpu().numpy(), b.cpu().numpy(), s.cpu().numpy().view(np.uint32))
    return out


@pytest.mark.parametrize("kw", [dict(M=301, K=16, D=45), dict(M=40, K=5, D=20, covariance="density", ragged=True)])
def test_static_scratch_entry_point_matches_existing_twice_in_a_row(eng, kw):
    ms = synth.make_model(seed=11, **kw)
    tab, _ = oracle_quant_tables(ms)
    eng.upload_quant(**tab)
    M = ms.M
    # two calls in a row, the second shorter than the first: the shared array still holds the first call's tail
    for T, seed in ((777, 1), (130, 2)):
        x = synth.make_features(ms, T, seed=seed)
        r = _both(eng, x, T, M)
        for a, b, what in zip(r["existing"], r["static"], ("q", "best", "score bits")):
            assert np.array_equal(a, b), f"{what} differ at T={T}"  # integer / bit-exact: the tests' bar for this scorer


def test_static_scratch_entry_point_rejects_sizes_beyond_its_limit(eng):
    ms = synth.make_model(M=8, K=4, D=45, seed=3)
    tab, _ = oracle_quant_tables(ms)
    eng.upload_quant(**tab)
    limit = capi.QUANT_STATIC_SCRATCH_BYTES // (eng.qC * eng.qDpad)
    with pytest.raises(capi.GmmError, match="exceed the static scratch"):
        eng.quant_score_device_static(limit + 1, 1, None, None, None)  # refused before anything is launched
