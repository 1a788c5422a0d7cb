"""CPU, world_size 2, gloo: the multi-GPU host logic -- segment ownership, mixture-shard tables and the all-gather
assembly of per-shard score blocks.  The per-shard scores come from the oracle here (test infrastructure); on GPUs
they come from the engine (tests/test_gpu_gather.py, profiles/bench_sharded.py, and bench.py's
extra.cfg5_mixture_sharded at N > 1)."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rasr_b200 import sharded, synth, tables


def test_segment_sharding_is_a_partition():
    for world in (1, 2, 4, 8):
        seen = sorted(s for r in range(world) for s in sharded.segments_of_rank(37, r, world))
        assert seen == list(range(37))
        assert all(sharded.segment_owner(s, world) == r for r in range(world)
                   for s in sharded.segments_of_rank(37, r, world))


def test_mixture_ranges_and_index_map():
    for M, world in ((4501, 8), (7, 2), (65536, 8), (5, 8)):
        mp_ = sharded.shard_width(M, world)
        cover = []
        for r in range(world):
            a, b = sharded.mixture_range(M, r, world)
            cover += list(range(a, b))
            for m in range(a, b):
                assert sharded.gathered_index(m, M, world) == (r, m - r * mp_)
        assert cover == list(range(M))


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle as O
        from rasr_b200.gms import MixtureSetData
        ms = synth.make_model(M=23, K=5, D=9, ragged=True, seed=4)
        x = synth.make_features(ms, 11)
        tab = tables.float_tables(ms)
        M = ms.M
        m0, m1 = sharded.mixture_range(M, rank, world)
        sub = sharded.shard_float_tables(tab, m0, m1)
        # score the shard with the oracle on an equivalent sub-model (slots are already untied in `tab`)
        G = int(sub["offsets"][-1])
        sub_ms = MixtureSetData(ms.D, sub["offsets"], np.arange(G, dtype=np.uint32),
                                ms.mix_logw[int(tab["offsets"][m0]):int(tab["offsets"][m1])],
                                np.arange(G, dtype=np.uint32), sub["cov_idx"], sub["mean"], ms.variances)
        s_loc, _ = O.score_max(sub_ms, x)
        local = torch.zeros((x.shape[0], sharded.shard_width(M, world)), dtype=torch.float32)
        local[:, : m1 - m0] = torch.from_numpy(s_loc)
        gathered = sharded.all_gather_scores(local, dist, world)
        full = sharded.to_frame_major(gathered, M).numpy()
        ref, _ = O.score_max(ms, x)
        ok = np.array_equal(full.view(np.uint32), ref.view(np.uint32))
        r_, c_ = sharded.gathered_index(M - 1, M, world)
        ok = ok and gathered[r_, 3, c_].item() == ref[3, M - 1]
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_mixture_sharded_all_gather_world2():
    from oracle import oracle as O
    O.build()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    res = dict(q.get(timeout=5) for _ in range(2))
    assert res == {0: True, 1: True}


def _train_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle as O
        from rasr_b200.align import AlignmentAutomata, linear_hmm
        ms = synth.make_model(M=31, K=4, D=8, seed=6, covariance="mixture")
        rng = np.random.default_rng(0)           # every rank builds the same corpus, then takes its segments
        n_seg = 7
        fo = np.cumsum([0] + [int(n) for n in rng.integers(20, 60, n_seg)]).astype(np.uint32)
        x = synth.make_features(ms, int(fo[-1]), seed=9)
        segs = [linear_hmm(rng.integers(0, ms.M, 6)) for _ in range(n_seg)]
        tab = O.float_tables(ms)
        scores, _ = O.score_max(ms, x, tables=tab)

        def e_step(seg_ids):
            acc = None
            for i in seg_ids:                    # align the segment, accumulate its frames on its path
                au = AlignmentAutomata.from_arc_lists([segs[i]])
                rows = slice(int(fo[i]), int(fo[i + 1]))
                em, _, _ = O.viterbi_align([0, rows.stop - rows.start], scores[rows], au)
                a = O.accumulate(ms, x[rows], em, tables=tab)
                acc = a if acc is None else {k: acc[k] + a[k] for k in acc if k != "best"}
            return acc

        mine = e_step(sharded.segments_of_rank(n_seg, rank, world))
        block = torch.from_numpy(sharded.pack_accumulators(mine))
        sharded.all_reduce_block(block, dist)
        got = sharded.unpack_accumulators(block.numpy(), ms.n_means, ms.n_cov, ms.D, ms.G)
        want = e_step(range(n_seg))
        ok = np.array_equal(got["mix_weight"], want["mix_weight"]) and got["mix_weight"].sum() == fo[-1]
        for k in sharded.ACC_FIELDS:
            ok = ok and np.allclose(got[k], want[k], rtol=1e-12, atol=1e-12)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_training_pass_segment_sharded_all_reduce_world2():
    """Segments i mod N per rank, alignment + E-step locally, one all-reduce of the packed estimators."""
    from oracle import oracle as O
    O.build()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_train_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    res = dict(q.get(timeout=5) for _ in range(2))
    assert res == {0: True, 1: True}
