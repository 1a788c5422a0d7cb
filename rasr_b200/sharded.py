"""Multi-GPU plumbing: segment sharding (no collective) and the mixture-sharded variant (one all-gather).

* Segment sharding is the reference's own scale-out scheme -- independent processes over corpus partitions
  (/root/reference/src/Bliss/CorpusDescription.cc:168-184,475-490): segment i belongs to rank i mod N, every rank
  holds the full model, nothing is exchanged.
* Mixture sharding (BASELINE.json configs[4], multi-million-Gaussian models): rank r holds mixtures
  [r*Mp, (r+1)*Mp) with Mp = ceil(M / N); every rank scores the SAME frames against its shard, then one
  ``all_gather`` (NCCL over NVLink on GPUs; gloo in the CPU tests) assembles the rank-major block
  ``[N][T][Mp]``.  ``gathered_index(m)`` maps an emission index to (rank, column); ``to_frame_major`` restores
  ``[T][M]`` when a consumer wants it.

* Training pass (alignment + E-step): segment sharded like scoring; the one real exchange is the sum of the
  ranks' estimators at the end of a pass -- the reference combines the accumulator files of its parallel jobs
  (AbstractMixtureSetEstimator::accumulate(const AbstractMixtureSetEstimator&),
  /root/reference/src/Mm/AbstractMixtureSetEstimator.cc:149-169); here it is one ``all_reduce`` over the packed
  f64 block on the device (``all_reduce_accumulators``).

* Push variant of mixture sharding (``PushMixtureShardedScorer``): no collective call at all -- each rank pushes its
  finished column ranges into every rank's frame-major block ``[T][N*Mp]`` through peer-mapped memory while it
  scores the next range, and the ranks synchronise through flag words on the device (include/gmm_b200.h
  gmm_score_device_gather); torch.distributed only exchanges the IPC handles.

torch.distributed is plumbing only; the scores come from the C-ABI engine (rasr_b200.capi).
"""
from __future__ import annotations

import numpy as np


# ---- segment sharding ------------------------------------------------------------------------------------
def segment_owner(segment: int, world: int) -> int:
    return segment % world


def segments_of_rank(n_segments: int, rank: int, world: int):
    return list(range(rank, n_segments, world))


# ---- mixture sharding ------------------------------------------------------------------------------------
def shard_width(M: int, world: int) -> int:
    return (M + world - 1) // world


def mixture_range(M: int, rank: int, world: int):
    mp = shard_width(M, world)
    return min(M, rank * mp), min(M, (rank + 1) * mp)


def gathered_index(m: int, M: int, world: int):
    mp = shard_width(M, world)
    return m // mp, m % mp


def shard_float_tables(tab: dict, m0: int, m1: int) -> dict:
    """Sub-model holding mixtures [m0, m1) of a float table dict (rasr_b200.tables.float_tables)."""
    offs = np.asarray(tab["offsets"], dtype=np.int64)
    g0, g1 = int(offs[m0]), int(offs[m1])
    out = dict(tab)
    out["offsets"] = (offs[m0:m1 + 1] - g0).astype(np.uint32)
    for k in ("mean", "cov_idx", "w2"):
        out[k] = np.ascontiguousarray(tab[k][g0:g1])
    return out


def shard_quant_tables(tab: dict, m0: int, m1: int) -> dict:
    offs = np.asarray(tab["offsets"], dtype=np.int64)
    g0, g1 = int(offs[m0]), int(offs[m1])
    out = dict(tab)
    out["offsets"] = (offs[m0:m1 + 1] - g0).astype(np.uint32)
    for k in ("qmean", "cov_idx", "qconst"):
        out[k] = np.ascontiguousarray(tab[k][g0:g1])
    return out


def all_gather_scores(local, dist, world: int):
    """local: torch tensor [T, Mp] (same shape on every rank) -> [world, T, Mp] on every rank."""
    import torch
    T = local.shape[0]
    out = torch.empty((world * T,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local.contiguous())  # rank-major concatenation along dim 0
    return out.view((world, T) + tuple(local.shape[1:]))


def to_frame_major(gathered, M: int):
    """[N][T][Mp] -> [T][M] (drops the padding columns of the last shard)."""
    n, T, mp = gathered.shape
    return gathered.permute(1, 0, 2).reshape(T, n * mp)[:, :M]


class MixtureShardedScorer:
    """One rank of the mixture-sharded float scorer: its own engine over mixtures [m0, m1)."""

    def __init__(self, float_tables: dict, rank: int, world: int, dist, device: int, mode=0, kernel=0,
                 total_mixtures: int | None = None):
        """float_tables: the whole model (sharded here), or -- with total_mixtures = M given -- this rank's shard only
        (mixtures [m0, m1) of an M-mixture model: a rank of a large model need not hold the other shards' tables)."""
        import torch
        from . import capi
        self.M = int(len(float_tables["offsets"]) - 1) if total_mixtures is None else int(total_mixtures)
        self.rank, self.world, self.dist = rank, world, dist
        self.Mp = shard_width(self.M, world)
        self.m0, self.m1 = mixture_range(self.M, rank, world)
        sub = shard_float_tables(float_tables, self.m0, self.m1) if total_mixtures is None else float_tables
        if int(len(sub["offsets"]) - 1) != self.m1 - self.m0:
            raise ValueError("the shard's tables do not hold this rank's mixtures")
        self.engine = capi.Engine(device)
        self.engine.upload_float(sub["offsets"], sub["mean"], sub["isr"], sub["cov_idx"], sub["w2"], sub["log_norm"],
                                 mode=mode, kernel=kernel)
        self._torch = torch
        self._local = None
        self._bound = False

    def bind_stream(self, stream):
        self.engine.set_stream(stream.cuda_stream)
        self._bound = True

    def score(self, d_x):
        """d_x: CUDA tensor [T, D] (identical on every rank) -> gathered [N][T][Mp] CUDA tensor."""
        torch = self._torch
        T = d_x.shape[0]
        if self._local is None or self._local.shape[0] != T:
            self._local = torch.zeros((T, self.Mp), dtype=torch.float32, device=d_x.device)
        if self.m1 > self.m0:
            self.engine.score_device(T, d_x, self._local, None, ld=self.Mp)
        if not self._bound:
            self.engine.sync()  # the engine's own stream is not the one the collective runs on
        return all_gather_scores(self._local, self.dist, self.world)


class _DeviceBlock:
    """A device allocation of the engine seen as a CUDA array (torch.as_tensor takes it without a copy)."""

    def __init__(self, ptr: int, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False), "version": 3}


class PushMixtureShardedScorer:
    """Mixture-sharded float scorer whose exchange step runs inside the scoring call, without NCCL and without a
    host barrier: every rank scores its shard into its own frame-major block [T][N*Mp], one range of B tiles at a
    time, and a push kernel on a second stream copies each finished column range into the peers' blocks (mapped
    through CUDA IPC; full-line stores over NVLink / NVSwitch) while the tensor kernel scores the next range.  The
    ranks synchronise on the device through flag words behind the score rows (include/gmm_b200.h
    gmm_score_device_gather); torch.distributed only carries the IPC handles at set-up."""

    FLAG_WORDS = 32

    def __init__(self, float_tables: dict, rank: int, world: int, dist, device: int, max_frames: int, mode=0,
                 chunks: int = 4, row_pad: int = 0, total_mixtures: int | None = None, n_blocks: int = 1):
        import torch
        from . import capi
        if world > 8:
            raise ValueError("at most 8 ranks")
        self.M = int(len(float_tables["offsets"]) - 1) if total_mixtures is None else int(total_mixtures)
        self.rank, self.world, self.dist = rank, world, dist
        self.Mp = shard_width(self.M, world)
        self.ld = self.Mp * world + int(row_pad)  # row_pad floats behind every row (a row stride off the power of two)
        self.m0, self.m1 = mixture_range(self.M, rank, world)
        if self.m1 <= self.m0:
            raise ValueError("a rank without mixtures: use fewer ranks")
        sub = shard_float_tables(float_tables, self.m0, self.m1) if total_mixtures is None else float_tables
        if int(len(sub["offsets"]) - 1) != self.m1 - self.m0:
            raise ValueError("the shard's tables do not hold this rank's mixtures")
        self.engine = capi.Engine(device)
        self.engine.upload_float(sub["offsets"], sub["mean"], sub["isr"], sub["cov_idx"], sub["w2"], sub["log_norm"],
                                 mode=mode, kernel=capi.KERNEL_TENSOR)
        if not self.engine.float_accuracy_ok:
            # the automatic kernel choice would have sent this shard to the direct kernel (means further out than
            # 8 sigma rms: the expanded form loses the 1e-4 relative bar); there is no direct variant of the push
            self.engine.close()
            raise ValueError("this shard's means are too far out for the tensor kernel's accuracy bar; use "
                             "MixtureShardedScorer (direct kernel + all-gather) for this model")
        self.max_frames, self.chunks = int(max_frames), int(chunks)
        self._consumer = None
        self.flags_offset = (self.max_frames * self.ld * 4 + 127) // 128 * 128
        # n_blocks = 2: two block sets used in turn by begin() / finish(), so that a call's last push runs under the
        # next call's scoring (include/gmm_b200.h gmm_score_device_gather_begin)
        dev = f"cuda:{device}"
        self._owns, self._all_bases, self.blocks = [], [], []
        for _ in range(max(1, int(n_blocks))):
            own, handle = self.engine.peer_alloc(self.flags_offset + self.FLAG_WORDS * 4)
            handles = [None] * world
            dist.all_gather_object(handles, handle)
            self._owns.append(own)
            self._all_bases.append([own if r == rank else self.engine.peer_open(handles[r]) for r in range(world)])
            self.blocks.append(torch.as_tensor(_DeviceBlock(own, (self.max_frames, self.ld)), device=dev))
        self._own, self._bases, self.block = self._owns[0], self._all_bases[0], self.blocks[0]
        self.flags = torch.as_tensor(_DeviceBlock(self._own + self.flags_offset, (self.FLAG_WORDS,)),
                                     device=dev).view(torch.int32)
        torch.cuda.synchronize()
        dist.barrier()

    def bind_stream(self, stream):
        """Run the engine on the torch stream whose work consumes the returned block (gmm_set_stream).  The READY
        signal of the next score() is then ordered behind every reader of the previous result, and readers queued
        after score() behind the wait for the peers' columns.  Without a bound stream the engine works on a stream of
        its own and the caller orders its consumers itself: engine.sync() after score(), and the consumers'
        stream synchronised before the next score()."""
        self.engine.set_stream(stream.cuda_stream)
        self._consumer = stream

    def score(self, d_x):
        """d_x: CUDA tensor [T, D] (identical on every rank) -> frame-major [T][N*Mp] view of this rank's block,
        complete for work queued on the engine's stream after the call (see bind_stream).  The view is
        peer-writable memory: the next score() lets the peers overwrite it."""
        T = d_x.shape[0]
        if T > self.max_frames:
            raise ValueError("more frames than the gathered block holds")
        self.engine.score_device_gather(T, d_x, self.rank, self._bases, self.ld, self.rank * self.Mp,
                                        self.flags_offset, self.chunks)
        return self.block[:T]

    def begin(self, d_x, k: int = 0):
        """Start a call on block set k (scoring + pushes) without joining it; finish(k) joins.  Used in turn --
        begin(x0, 0); begin(x1, 1); finish(0); begin(x2, 0); finish(1); ... -- a call's exchange runs under the next
        call's scoring."""
        T = d_x.shape[0]
        if T > self.max_frames:
            raise ValueError("more frames than the gathered block holds")
        self.engine.score_device_gather_begin(T, d_x, self.rank, self._all_bases[k], self.ld, self.rank * self.Mp,
                                              self.flags_offset, self.chunks)
        return T

    def finish(self, k: int = 0, T: int | None = None):
        """Join the call started on block set k -> its frame-major view (complete for work queued on the engine's
        stream after this)."""
        self.engine.gather_finish(self._owns[k], self.flags_offset)
        return self.blocks[k] if T is None else self.blocks[k][:T]

    def timed_out(self) -> bool:
        """A peer did not show up within the device-side limit (call after a sync).  The other ranks' columns of
        that call read +inf, and scoring calls raise until reset()."""
        return self.engine.gather_status() != 0

    def reset(self):
        """After a time-out, on every rank: clears the flag words and the condition."""
        self.engine.sync()
        self.dist.barrier()
        self.engine.gather_reset()
        self.dist.barrier()

    def close(self):
        self.engine.sync()
        self.dist.barrier()
        for bases in self._all_bases:
            for r, b in enumerate(bases):
                if r != self.rank:
                    self.engine.peer_close(b)
        self.dist.barrier()
        self.blocks, self.block = [], None
        for own in self._owns:
            self.engine.peer_free(own)
        self.engine.close()


# ---- training pass: combine the ranks' estimators ----------------------------------------------------------
ACC_FIELDS = ("mean_sum", "mean_weight", "cov_sum", "cov_weight", "mix_weight")


def pack_accumulators(acc: dict) -> np.ndarray:
    """dict of Engine.read_accumulators / oracle.accumulate -> the block layout of gmm_accumulator_export_device."""
    return np.concatenate([np.asarray(acc[k], np.float64).ravel() for k in ACC_FIELDS])


def unpack_accumulators(block, n_means: int, n_cov: int, D: int, G: int) -> dict:
    block = np.asarray(block, np.float64)
    sizes = (n_means * D, n_means, n_cov * D, n_cov, G)
    assert block.shape == (sum(sizes),)
    parts = np.split(block, np.cumsum(sizes)[:-1])
    return dict(mean_sum=parts[0].reshape(n_means, D), mean_weight=parts[1], cov_sum=parts[2].reshape(n_cov, D),
                cov_weight=parts[3], mix_weight=parts[4])


def all_reduce_block(block, dist):
    """Sum the packed estimator block over all ranks, in place (NCCL over NVLink on GPUs, gloo in the CPU tests)."""
    dist.all_reduce(block, op=dist.ReduceOp.SUM)
    return block


def all_reduce_accumulators(engine, dist, stream=None):
    """Every rank ends up with the sum of all ranks' estimators, without leaving the device."""
    import torch
    n = engine.accumulator_size
    block = torch.empty(n, dtype=torch.float64, device="cuda")
    engine.export_accumulators(block)
    if stream is None:
        engine.sync()  # the engine's own stream is not the one the collective runs on
    all_reduce_block(block, dist)
    engine.import_accumulators(block)
    if stream is None:
        torch.cuda.current_stream().synchronize()
        engine.sync()
    return block
