"""Mixture-sharded variant (BASELINE.json configs[4]) under torchrun: every rank scores the same T frames against its
M/N mixtures, then one NCCL all-gather of the [T][M/N] blocks.  Prints one JSON line (rank 0).
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 profiles/bench_sharded.py [M K T steps]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from rasr_b200 import sharded, synth, tables

M = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
K = int(sys.argv[2]) if len(sys.argv) > 2 else 64
T = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 20
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
# every rank builds only its own shard of the seeded model (same generator stream -> identical global model)
mp_ = sharded.shard_width(M, world)
m0, m1 = sharded.mixture_range(M, rank, world)
full = synth.make_model(M, K, 45, seed=1001)
tab = tables.float_tables(full)
x = synth.make_features(full, T)
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
sc = sharded.MixtureShardedScorer(tab, rank, world, dist, local)
sc.bind_stream(stream)
dx = torch.from_numpy(x).cuda()
for _ in range(3):
    g = sc.score(dx)
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
e0.record()
for _ in range(steps):
    g = sc.score(dx)
e1.record()
for _ in range(steps):  # kernel only, for the split
    sc.engine.score_device(T, dx, sc._local, None, ld=sc.Mp)
e2.record()
torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1), e1.elapsed_time(e2)], device="cuda", dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
# correctness: gathered block equals a single-engine score of a sample of mixtures
ok = True
ms_total, ms_kernel = t[0].item() / steps, t[1].item() / steps
if rank == 0:
    fm = sharded.to_frame_major(g, M)
    from rasr_b200 import capi
    probe = sharded.shard_float_tables(tab, M - 7, M)
    e = capi.Engine(local)
    e.upload_float(probe["offsets"], probe["mean"], probe["isr"], probe["cov_idx"], probe["w2"], probe["log_norm"])
    s, _ = e.score_host(x, want_best=False)
    ok = bool(np.allclose(fm[:, M - 7:].cpu().numpy(), s, rtol=1e-4, atol=1e-4))
    print(json.dumps({"workload": f"mixture-sharded {M}x{K}x45, T={T}", "n_gpus": world, "ms_per_step": ms_total,
                      "ms_kernel": ms_kernel, "ms_all_gather": ms_total - ms_kernel,
                      "frames_per_s": T / (ms_total * 1e-3), "gather_bytes_per_rank": T * mp_ * 4,
                      "gathered_bytes": world * T * mp_ * 4,
                      "allgather_busbw_GBs": (world - 1) * T * mp_ * 4 / max(1e-9, (ms_total - ms_kernel) * 1e-3) / 1e9,
                      "gaussians_per_rank": (m1 - m0) * K, "check": ok}))
# ---- push variant: column ranges pushed into the peers' blocks while the next range is scored, flags on device ----
def timed_loop(fn):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier()
    ev0.record(stream)
    for _ in range(steps):
        fn()
    ev1.record(stream)
    torch.cuda.synchronize()
    tt = torch.tensor([ev0.elapsed_time(ev1) / steps], device="cuda", dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    return tt.item()

ref_fm = sharded.to_frame_major(sc.score(dx), M).clone()
torch.cuda.synchronize()
out = {"workload": f"mixture-sharded {M}x{K}x45, T={T}, push variant (no NCCL call)", "n_gpus": world,
       "ms_per_step_all_gather": ms_total}
for chunks in (1, 2, 4, 8):
    ps = sharded.PushMixtureShardedScorer(tab, rank, world, dist, local, max_frames=T, chunks=chunks,
                                          row_pad=int(os.environ.get("PUSH_ROW_PAD", "0")))
    ps.engine.set_stream(stream.cuda_stream)
    for _ in range(3):
        fm = ps.score(dx)
    ms = timed_loop(lambda: ps.score(dx))
    fm = ps.score(dx)
    torch.cuda.synchronize()
    same = torch.tensor([1.0 if torch.equal(fm[:, :M], ref_fm) and not ps.timed_out() else 0.0], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    out[f"ms_per_step_push_{chunks}_ranges"] = ms
    out[f"bit_identical_{chunks}_ranges"] = bool(same.item() == 1.0)
    ps.close()
# ---- scores consumed on the HOST: every rank copies its own [T][Mp] columns to its own pinned block
# over its own PCIe link, against exchanging first and copying the whole [T][N*Mp] block from one GPU ----
own_host = torch.empty((T, sc.Mp), dtype=torch.float32, pin_memory=True)
def score_and_copy_own():
    sc.engine.score_device(T, dx, sc._local, None, ld=sc.Mp)
    own_host.copy_(sc._local.view(T, sc.Mp), non_blocking=True)
for _ in range(2):
    score_and_copy_own()
out["ms_per_step_host_consumed_own_columns_d2h"] = timed_loop(score_and_copy_own)
ps = sharded.PushMixtureShardedScorer(tab, rank, world, dist, local, max_frames=T, chunks=4)
ps.engine.set_stream(stream.cuda_stream)
all_host = torch.empty((T, ps.ld), dtype=torch.float32, pin_memory=True)
def exchange_and_copy_all():
    all_host.copy_(ps.score(dx), non_blocking=True)
for _ in range(2):
    exchange_and_copy_all()
out["ms_per_step_host_consumed_exchange_then_d2h_all"] = timed_loop(exchange_and_copy_all)
ps.close()
best = min(v for k, v in out.items() if k.startswith("ms_per_step_push"))
out["frames_per_s_push"] = T / (best * 1e-3)
out["speedup_over_all_gather"] = ms_total / best
if rank == 0:
    print(json.dumps(out))
dist.barrier()
dist.destroy_process_group()
