"""GPU: the mixture-sharded scoring call with its exchange step (gmm_score_device_gather).  On one GPU it runs with
world = 1 (the shard scored one range of B tiles at a time into a block the library allocated) and must give the very
bits of gmm_score_device; with two GPUs visible, two processes push their columns into each other's blocks and every
rank must hold the frame-major scores of the whole model (profiles/bench_sharded.py measures the same path)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from rasr_b200 import capi, sharded, synth, tables

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def upload(eng, ms, mode=capi.MODE_MAX):
    t = tables.float_tables(ms)
    eng.upload_float(t["offsets"], t["mean"], t["isr"], t["cov_idx"], t["w2"], t["log_norm"], mode=mode,
                     kernel=capi.KERNEL_TENSOR)


# T / chunks >= 1024 frames: the call is cut into frame ranges, else into column ranges (gmm_gather.cu)
@pytest.mark.parametrize("M,K,D,ragged,mode,chunks,T", [(4501, 16, 45, False, capi.MODE_MAX, 4, 1000), (4501, 16, 45, False, capi.MODE_MAX, 8, 1000),
                                                        (300, 8, 20, False, capi.MODE_SUM, 3, 1000), (97, 5, 12, True, capi.MODE_MAX, 4, 1000),
                                                        (33, 64, 45, False, capi.MODE_MAX, 1, 1000), (300, 8, 20, False, capi.MODE_MAX, 2, 2500),
                                                        (97, 5, 12, True, capi.MODE_SUM, 3, 3333)])
def test_ranges_equal_one_pass(M, K, D, ragged, mode, chunks, T):
    ms = synth.make_model(M=M, K=K, D=D, seed=M + K, ragged=ragged)
    x = synth.make_features(ms, T, seed=5)
    eng = capi.Engine(0)
    try:
        upload(eng, ms, mode)
        dx = torch.from_numpy(x).cuda()
        full = torch.empty((T, M), device="cuda")
        eng.score_device(T, dx, full)
        col, ld = 8, M + 24  # the shard's columns sit inside a wider row
        flags_offset = (T * ld * 4 + 127) // 128 * 128
        base, _ = eng.peer_alloc(flags_offset + 128)
        block = torch.as_tensor(sharded._DeviceBlock(base, (T, ld)), device="cuda:0")
        n0 = eng.launch_count
        eng.score_device_gather(T, dx, 0, [base], ld, col, flags_offset, chunks)
        eng.sync()
        assert eng.launch_count > n0
        got = block[:, col:col + M]
        assert torch.equal(got.view(torch.int32), full.view(torch.int32))
        assert float(block[:, :col].abs().max()) == 0.0 and float(block[:, col + M:].abs().max()) == 0.0
        del block, got
        eng.peer_free(base)
    finally:
        eng.close()


def test_aligner_reads_the_gathered_block_in_place():
    """Column m of the block is mixture m: the device-resident aligner takes the block (pointer + ld) like any score
    matrix and finds the path it finds over gmm_score_device's scores."""
    from rasr_b200.align import AlignmentAutomata, linear_hmm
    ms = synth.make_model(M=600, K=8, D=24, seed=21)
    T = 700
    x = synth.make_features(ms, T, seed=4)
    rng = np.random.default_rng(5)
    fo = np.array([0, 300, 700], np.uint32)
    au = AlignmentAutomata.from_arc_lists([linear_hmm(rng.integers(0, ms.M, n)) for n in (70, 120)])
    eng = capi.Engine(0)
    try:
        upload(eng, ms)
        dx = torch.from_numpy(x).cuda()
        full = torch.empty((T, ms.M), device="cuda")
        eng.score_device(T, dx, full)
        want = eng.align_device(fo, full, au, ld=ms.M)
        ld = ms.M + 40
        flags_offset = (T * ld * 4 + 127) // 128 * 128
        base, _ = eng.peer_alloc(flags_offset + 128)
        eng.score_device_gather(T, dx, 0, [base], ld, 0, flags_offset, 4)
        got = eng.align_device(fo, base, au, ld=ld)
        for g, w in zip(got, want):
            assert np.array_equal(g.view(np.uint32), w.view(np.uint32))
        eng.peer_free(base)
    finally:
        eng.close()


def test_gather_rejects_bad_arguments():
    ms = synth.make_model(M=40, K=4, D=10, seed=1)
    eng = capi.Engine(0)
    try:
        upload(eng, ms)
        dx = torch.zeros((8, 10), device="cuda")
        base, _ = eng.peer_alloc(8 * 64 * 4 + 128)
        with pytest.raises(capi.GmmError):  # the shard's columns do not fit in ld
            eng.score_device_gather(8, dx, 0, [base], 32, 0, 8 * 32 * 4, 1)
        with pytest.raises(capi.GmmError):  # flag words inside the score rows
            eng.score_device_gather(8, dx, 0, [base], 64, 0, 128, 1)
        with pytest.raises(capi.GmmError):  # rank outside the world
            eng.score_device_gather(8, dx, 1, [base], 64, 0, 8 * 64 * 4, 1)
        eng.peer_free(base)
    finally:
        eng.close()


def _shard_engines(ms, world, mode=capi.MODE_MAX):
    tab = tables.float_tables(ms)
    M = len(tab["offsets"]) - 1
    engines = []
    for r in range(world):
        m0, m1 = sharded.mixture_range(M, r, world)
        sub = sharded.shard_float_tables(tab, m0, m1)
        e = capi.Engine(0)
        e.upload_float(sub["offsets"], sub["mean"], sub["isr"], sub["cov_idx"], sub["w2"], sub["log_norm"], mode=mode,
                       kernel=capi.KERNEL_TENSOR)
        engines.append((e, m0, m1))
    return engines, M


# two handles only: with more, their streams share hardware queues and a spinning wait kernel of one handle sits in
# front of another handle's scoring kernels (again a one-GPU artefact)
@pytest.mark.parametrize("world,chunks,T", [(2, 4, 900), (2, 2, 2304), (2, 1, 130)])
def test_ranks_on_one_gpu_exchange_their_columns(world, chunks, T):
    """Push / DONE / wait between `world` handles of ONE process on one GPU: every handle has its own streams and its
    own block, the blocks are plain device pointers to each other.  The READY words are preset: on one GPU a push
    kernel spinning for a peer's READY keeps the SMs the next scoring kernel wants, and the peer's signal kernel
    queues behind that one -- a scheduling dead end that separate GPUs do not have (the READY handshake is covered
    by the two-GPU test below and by profiles/bench_sharded.py)."""
    ms = synth.make_model(M=1000, K=8, D=24, seed=77)
    x = synth.make_features(ms, T, seed=9)
    engines, M = _shard_engines(ms, world)
    Mp = sharded.shard_width(M, world)
    ld = Mp * world
    flags_offset = (T * ld * 4 + 127) // 128 * 128
    try:
        dx = torch.from_numpy(x).cuda()
        want = torch.zeros((T, ld), device="cuda")
        for r, (e, m0, m1) in enumerate(engines):
            own = torch.empty((T, m1 - m0), device="cuda")
            e.score_device(T, dx, own)
            e.sync()
            want[:, r * Mp:r * Mp + (m1 - m0)] = own
        bases = [e.peer_alloc(flags_offset + 128)[0] for e, _, _ in engines]
        blocks = [torch.as_tensor(sharded._DeviceBlock(b, (T, ld)), device="cuda:0") for b in bases]
        flags = [torch.as_tensor(sharded._DeviceBlock(b + flags_offset, (32,)), device="cuda:0").view(torch.int32) for b in bases]
        for f in flags:
            f[:8] = 1 << 20
        torch.cuda.synchronize()
        for rep in range(3):
            for r, (e, _, _) in enumerate(engines):
                e.score_device_gather(T, dx, r, bases, ld, r * Mp, flags_offset, chunks)
            for e, _, _ in engines:
                e.sync()
            for r in range(world):
                assert int(flags[r][17]) == 0, "time-out"
                assert torch.equal(blocks[r].view(torch.int32), want.view(torch.int32)), (rep, r)
                blocks[r].zero_()
            torch.cuda.synchronize()
        del blocks, flags
        for (e, _, _), b in zip(engines, bases):
            e.peer_free(b)
    finally:
        for e, _, _ in engines:
            e.close()


def test_begin_and_finish_over_two_block_sets_in_turn():
    """gmm_score_device_gather_begin / gmm_gather_finish: two handles of one process, two block sets each, calls started
    on one set while the other is joined -- every joined block equals the scores of the frames its call was given, and a
    begin on a block whose call was never joined joins it first."""
    ms = synth.make_model(M=1000, K=8, D=24, seed=78)
    world, T = 2, 700
    engines, M = _shard_engines(ms, world)
    Mp = sharded.shard_width(M, world)
    ld = Mp * world
    flags_offset = (T * ld * 4 + 127) // 128 * 128
    try:
        xs = [torch.from_numpy(synth.make_features(ms, T, seed=20 + i)).cuda() for i in range(3)]
        wants = []
        for dx in xs:
            want = torch.zeros((T, ld), device="cuda")
            for r, (e, m0, m1) in enumerate(engines):
                own = torch.empty((T, m1 - m0), device="cuda")
                e.score_device(T, dx, own)
                e.sync()
                want[:, r * Mp:r * Mp + (m1 - m0)] = own
            wants.append(want)
        sets = []  # [block set][rank] -> base pointer
        for _ in range(2):
            sets.append([e.peer_alloc(flags_offset + 128)[0] for e, _, _ in engines])
        views = [[torch.as_tensor(sharded._DeviceBlock(b, (T, ld)), device="cuda:0") for b in bases] for bases in sets]
        for bases in sets:
            for b in bases:
                f = torch.as_tensor(sharded._DeviceBlock(b + flags_offset, (32,)), device="cuda:0").view(torch.int32)
                f[:8] = 1 << 20  # READY preset: one-GPU artefact, see test_ranks_on_one_gpu_exchange_their_columns
        torch.cuda.synchronize()

        def begin(k, i):
            for r, (e, _, _) in enumerate(engines):
                e.score_device_gather_begin(T, xs[i], r, sets[k], ld, r * Mp, flags_offset, 2)

        def finish(k):
            for r, (e, _, _) in enumerate(engines):
                e.gather_finish(sets[k][r], flags_offset)
            for e, _, _ in engines:
                e.sync()

        begin(0, 0)
        begin(1, 1)
        finish(0)
        for r in range(world):
            assert torch.equal(views[0][r].view(torch.int32), wants[0].view(torch.int32))
        begin(0, 2)          # set 0 again while set 1 is still open
        finish(1)
        for r in range(world):
            assert torch.equal(views[1][r].view(torch.int32), wants[1].view(torch.int32))
        begin(1, 0)          # set 0's call (frames 2) is never joined explicitly ...
        begin(0, 1)          # ... the next begin on it joins it first
        finish(0)
        finish(1)
        for r in range(world):
            assert torch.equal(views[0][r].view(torch.int32), wants[1].view(torch.int32))
            assert torch.equal(views[1][r].view(torch.int32), wants[0].view(torch.int32))
            assert engines[r][0].gather_status() == 0
        del views
        for bases in sets:
            for (e, _, _), b in zip(engines, bases):
                e.peer_free(b)
    finally:
        for e, _, _ in engines:
            e.close()


def test_a_missing_peer_is_reported_poisoned_and_recoverable(monkeypatch):
    """A peer that never shows up: the wait ends, the caller can SEE it (gmm_gather_status, GMM_ERR_TIMEOUT from the
    next scoring / consuming call), a consumer already queued cannot read stale columns (+inf), and after
    gmm_gather_reset the handle works again."""
    monkeypatch.setenv("GMM_GATHER_TIMEOUT_MS", "150")
    ms = synth.make_model(M=64, K=4, D=10, seed=3)
    engines, M = _shard_engines(ms, 2)
    T, ld = 64, 64
    Mp = 32
    flags_offset = T * ld * 4
    try:
        dx = torch.from_numpy(synth.make_features(ms, T, seed=1)).cuda()
        e0 = engines[0][0]
        bases = [e0.peer_alloc(flags_offset + 128)[0] for _ in range(2)]
        blocks = [torch.as_tensor(sharded._DeviceBlock(b, (T, ld)), device="cuda:0") for b in bases]
        flags = [torch.as_tensor(sharded._DeviceBlock(b + flags_offset, (32,)), device="cuda:0").view(torch.int32) for b in bases]
        blocks[0].fill_(-7.0)  # "stale scores of an earlier call"
        torch.cuda.synchronize()
        assert e0.gather_status() == 0
        e0.score_device_gather(T, dx, 0, bases, ld, 0, flags_offset, 1)  # rank 1 never calls: returns, no error yet
        e0.sync()
        assert int(flags[0][17]) == 1          # marked on the waiting rank ...
        assert int(flags[1][8 + 0]) == 0       # ... which did not announce its columns as complete
        assert e0.gather_status() == 1         # ... and says so to the host: call number 1
        own = torch.empty((T, Mp), device="cuda")
        e0.score_device(T, dx, own)
        e0.sync()
        assert torch.equal(blocks[0][:, :Mp], own)                  # own columns are the real scores
        assert bool(torch.isinf(blocks[0][:, Mp:]).all())           # the peer's columns: +inf, not the stale -7
        with pytest.raises(capi.GmmError, match="timed out"):
            e0.score_device_gather(T, dx, 0, bases, ld, 0, flags_offset, 1)
        from rasr_b200.align import AlignmentAutomata, linear_hmm
        au = AlignmentAutomata.from_arc_lists([linear_hmm(np.arange(10) % Mp)])
        with pytest.raises(capi.GmmError, match="timed out"):  # a consumer of the block on this handle
            e0.align_device(np.array([0, T], np.uint32), bases[0], au, ld=ld)
        # a second handle that meets the marked block refuses too
        with pytest.raises(capi.GmmError, match="timed-out"):
            engines[1][0].score_device_gather(T, dx, 0, bases, ld, 0, flags_offset, 1)
        e0.gather_reset()
        assert e0.gather_status() == 0 and int(flags[0][17]) == 0
        flags[0][:8] = 1 << 20                 # play the peer: READY and DONE far ahead
        flags[0][8:16] = 1 << 20
        torch.cuda.synchronize()
        e0.score_device_gather(T, dx, 0, bases, ld, 0, flags_offset, 1)
        e0.sync()
        assert e0.gather_status() == 0
        assert torch.equal(blocks[0][:, :Mp], own)
        del flags, blocks
        for b in bases:
            e0.peer_free(b)
    finally:
        for e, _, _ in engines:
            e.close()


def test_a_new_handle_continues_the_blocks_call_numbering():
    """The call counter lives with the flag words: a fresh handle over used blocks does not start again at 1 (where
    every wait would be satisfied at once by the words of earlier calls)."""
    ms = synth.make_model(M=64, K=4, D=10, seed=3)
    T, ld, Mp = 64, 64, 32
    flags_offset = T * ld * 4
    engines, M = _shard_engines(ms, 2)
    try:
        dx = torch.from_numpy(synth.make_features(ms, T, seed=1)).cuda()
        e0 = engines[0][0]
        bases = [e0.peer_alloc(flags_offset + 128)[0] for _ in range(2)]
        flags = [torch.as_tensor(sharded._DeviceBlock(b + flags_offset, (32,)), device="cuda:0").view(torch.int32) for b in bases]
        for f in flags:
            f[:16] = 1 << 20
        torch.cuda.synchronize()
        for _ in range(3):
            e0.score_device_gather(T, dx, 0, bases, ld, 0, flags_offset, 1)
        e0.sync()
        assert int(flags[0][18]) == 3 and int(flags[1][8 + 0]) == 3
        e0.close()
        engines[0] = (None, 0, 0)
        # a new engine for the same rank over the same blocks: its first call is number 4
        fresh, _ = _shard_engines(ms, 2)
        e0b = fresh[0][0]
        fresh[1][0].close()
        try:
            e0b.score_device_gather(T, dx, 0, bases, ld, 0, flags_offset, 1)
            e0b.sync()
            assert int(flags[0][18]) == 4 and int(flags[1][8 + 0]) == 4
            del flags
            for b in bases:
                e0b.peer_free(b)
        finally:
            e0b.close()
    finally:
        for e, _, _ in engines:
            if e is not None:
                e.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_ranks_push_into_each_other():
    env = dict(os.environ, PYTHONPATH=ROOT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "profiles", "bench_sharded.py"), "2048", "16", "640", "3"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    import json
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    push = [l for l in lines if "push variant" in l["workload"]][0]
    for c in (1, 2, 4, 8):
        assert push[f"bit_identical_{c}_ranges"] is True
