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

    def __init__(self, float_tables: dict, rank: int, world: int, dist, device: int, mode=0, kernel=0):
        import torch
        from . import capi
        self.M = int(len(float_tables["offsets"]) - 1)
        self.rank, self.world, self.dist = rank, world, dist
        self.Mp = shard_width(self.M, world)
        self.m0, self.m1 = mixture_range(self.M, rank, world)
        sub = shard_float_tables(float_tables, self.m0, self.m1)
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
