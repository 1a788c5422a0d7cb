#!/usr/bin/env python
"""bench.py -- scored frames/s of the GMM emission-scoring hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch of synthetic frames: S utterance segments of 1000 frames of
45-dim features scored against the 4501 x 16-density mixture set (float maximum-approximation scorer, the
reference's `diagonal-maximum`), i.e. a slice of BASELINE.json configs[3] (corpus-scale, segment-sharded) on
the configs[0]/[1] model.  With N > 1 (torchrun, one rank per GPU) every rank scores its own segments
(segment i -> rank i mod N, no collective on the data path): weak scaling.

`value`   : whole-job frames/s, features and scores resident in HBM, CUDA-event timed, max over ranks.
`e2e`     : the same through the C-ABI host call (gmm_score_host): pinned host features in, host scores out,
            both copies inside the timed region.
`roofline`: dominant kernel, algorithmic flops 2*(2D+1)*G per frame / CUDA-event time,
            against MEASURED_PEAKS.json.
`cpu_baseline`: the reference's own scorer classes (oracle/_ref/ref_scorer, built from the unmodified reference
            sources) timed on this box's host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "scored frames/sec (45-dim, 4501x16 dens.)"
UNIT = "frames/s"
SEG_FRAMES = 1000
L2_BYTES = 126 << 20


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=float(d["hbm_gbs"]), tensor=float(d["bf16_tflops"]),
                    tensor_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, tensor=1590.0, tensor_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.thread = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for (ts, r) in self.rows if t0 - 0.05 <= ts <= t1 + 0.15 and len(r) >= 7] or \
               [r for (_, r) in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": float(rows[0][1]),
                "power_w_max": max(float(r[2]) for r in rows), "samples": len(rows), "reasons": reasons}


def bind_to_gpu_numa_node(index: int):
    """One process per GPU: pin this process to the CPUs next to its GPU (NVML's ideal affinity) so that page-locked
    buffers and the copy threads sit on the GPU's NUMA node; without it the 8-rank host-delivered rate is limited by
    cross-socket traffic.  Best effort: silently skipped when NVML is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(index))
        return True
    except Exception:
        return False


def workload(segments: int):
    from rasr_b200 import synth
    ms = synth.make_model(4501, 16, 45, seed=1001, covariance="pooled")
    return ms, segments * SEG_FRAMES


def segment_features(ms, seg_ids):
    """Features of whole segments, regenerated from seed 3000 + segment id."""
    from rasr_b200 import synth
    return np.concatenate([synth.make_features(ms, SEG_FRAMES, seed=3000 + int(s)) for s in seg_ids])


# -------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own CPU scorer on the host cores
# -------------------------------------------------------------------------------------------------
def cpu_reference_run(ms, frames: int, scorer: str, repeat: int, threads: int | None = None):
    """Times oracle/_ref/ref_scorer (kind "reference") or, if it was not built, the oracle port ("port")."""
    from oracle import oracle as O
    from rasr_b200 import synth
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    cores = cores if threads is None else min(cores, threads)
    x = synth.make_features(ms, frames, seed=3000)
    if O.ref_available():
        r = O.run_ref(ms, x, scorer, threads=cores, repeat=repeat, dump=False)
        fps = float(r["timing"]["frames_per_s"])
        kind = "reference"
    else:
        O.build()
        os.environ.setdefault("OMP_NUM_THREADS", str(cores))
        f = {"max": O.score_max, "sum": O.score_sum}[scorer]
        best = 1e30
        for _ in range(repeat):
            t0 = time.perf_counter()
            f(ms, x)
            best = min(best, time.perf_counter() - t0)
        fps, kind = frames / best, "port"
    return dict(value=fps, unit=UNIT, cores=cores, kind=kind, scorer=scorer,
                sample=f"{frames} frames of the same model/features, all {ms.M} mixtures per frame, "
                       f"one scorer instance per thread on {cores} threads, best of {repeat}")


def cpu_reference_legs(ms, cores: int, scorers=("max", "sum", "simd", "batchint")):
    """the reference's scorers on one thread and on all host cores, bounded samples (about a
    second each).  Only with the compiled reference (oracle/_ref/ref_scorer); reported beside the GPU numbers."""
    from oracle import oracle as O
    if not O.ref_available():
        return None
    out = {}
    per_thread = max(16, int(400e6 / max(1, ms.G) / 45))   # frames per thread: about 0.7 s of the float scorer
    for name in scorers:
        try:
            one = cpu_reference_run(ms, per_thread, name, 1, threads=1)
            allc = cpu_reference_run(ms, per_thread * cores, name, 1)
        except Exception as e:  # a scorer the reference refuses for this model (batch-int needs a pooled covariance)
            out[name] = {"unavailable": str(e)[:200]}
            continue
        out[name] = {"one_thread_frames_per_s": one["value"], "all_cores_frames_per_s": allc["value"],
                     "cores": allc["cores"], "frames_one_thread": per_thread, "frames_all_cores": per_thread * cores}
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    ms, _ = workload(1)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    frames = max(256, 48 * cores)  # about 1-2 s of CPU work per step
    for _ in range(args.warmup):
        cpu_reference_run(ms, frames, "max", 1)
    vals = []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        vals.append(cpu_reference_run(ms, frames, "max", 1))
    wall = time.perf_counter() - t0
    fps = statistics.median(v["value"] for v in vals)
    cb = dict(vals[0], value=fps)
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * frames / fps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "cfg4-slice: 4501x16x45 float diagonal-maximum, reference CPU scorer",
                       "frames_per_step": frames, "wall_s": wall,
                       "note": "a bounded sample of the GPU arm's workload: the CPU scorer evaluates frame by frame, "
                               "its frames/s does not depend on how many frames a step holds"},
            "cpu_baseline": cb, "gpu_launches": 0,
            "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


# -------------------------------------------------------------------------------------------------
# our arm
# -------------------------------------------------------------------------------------------------
def time_device(torch, fn, steps, warmup, dist=None):
    """W warm-up steps, then exactly K steps bracketed by barrier + synchronize; CUDA events on the current
    stream (the engine is bound to it); returns (seconds max over ranks, wall t0, wall t1)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.time()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    w1 = time.time()
    if dist is not None:
        dist.barrier()
    sec = e0.elapsed_time(e1) * 1e-3
    if dist is not None:
        t = torch.tensor([sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item())
    return sec, w0, w1


def ncu_traffic(summary_name):
    """dram read + write bytes per launch of the dominant kernel, from the committed ncu summary (profiles/)."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", summary_name)
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:
        for line in open(path):
            if line.startswith("traffic ="):
                tok = line.split(")")[-1].split()
                return float(tok[0]) * mult[tok[1]] + float(tok[3]) * mult[tok[4]]
    except (OSError, KeyError, ValueError, IndexError):
        pass
    return None


def float_roofline(ms, frames, kernel_sec, peaks, kernel_name, step_sec=None, traffic=None, sustained=False,
                   launches_per_step=1, split=3):
    """frames, kernel_sec: frames of ONE launch of the scoring kernel and its average duration (CUDA events around the
    launch; a step of `launches_per_step` launches: long batches go out in L2-sized launches).
    sustained: the kernel ran back to back for >= 100 ms (power-capped clocks) -> the sustained cuBLAS figure is
    the denominator; a kernel timed in a short burst is compared with the burst figure."""
    flops = 2.0 * (2 * ms.D + 1) * ms.G * frames  # algorithmic, SURVEY.md 8(d)
    ach = flops / kernel_sec / 1e12
    peak = peaks["tensor_sustained"] if sustained else peaks["tensor"]
    return {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
            "frac": ach / peak, "traffic": traffic, "kernel": kernel_name,
            "kernel_ms_per_launch": kernel_sec * 1e3,
            "launches_per_step": launches_per_step, "frames_per_launch": frames,
            "kernel_share_of_step": (kernel_sec * launches_per_step / step_sec) if step_sec else None,
            "frac_of_burst_peak": ach / peaks["tensor"],
            # MMA instruction times per tile in fp16-K16 equivalents over the algorithmic contraction length: `split`
            # products (3: fp16 hi.hi + lo.hi + hi.lo; 2: fp16 hi.hi + one e4m3 instruction for both corrections)
            "mma_products_per_algorithmic_product": split,
            "executed_over_algorithmic": float(split) * (16 * (((ms.D if ms.variances.shape[0] == 1 else 2 * ms.D) + 3 + 15) // 16))
                                         / (2 * ms.D + 1),
            "peak_source": f"{peaks['source']} bf16 dense {'sustained' if sustained else 'burst'} (MEASURED_PEAKS.json)",
            "algorithmic_flops_per_launch": flops,
            "gaussian_evals_per_s_kernel": ms.G * frames / kernel_sec}


def quant_roofline(ms, frames, kernel_sec, peaks, outputs=2, traffic=None, kernel_name="quant_score_kernel"):
    dpad = ((ms.D + 15) // 16) * 16
    nbytes = ms.G * (dpad + 4) + frames * dpad + frames * ms.M * 4 * outputs
    ach = nbytes / kernel_sec / 1e9
    return {"bound": "hbm", "achieved": ach, "peak": peaks["hbm"], "unit": "GB/s", "frac": ach / peaks["hbm"],
            "traffic": traffic, "kernel": kernel_name, "kernel_ms_per_launch": kernel_sec * 1e3,
            "outputs_written": outputs,
            "algorithmic_bytes_per_launch": nbytes,
            "peak_source": f"{peaks['source']} copy bandwidth (MEASURED_PEAKS.json)"}


def run_ours(args):
    import torch
    from rasr_b200 import capi, sharded, synth, tables
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        # NCCL's version banner / debug output goes to stderr: stdout carries the one JSON line only
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    peaks = measured_peaks()
    segs = args.segments
    ms, frames = workload(segs)
    tab = tables.float_tables(ms)
    all_cpus = os.sched_getaffinity(0)
    bind_to_gpu_numa_node(local)  # before any pinned allocation: host buffers land on the GPU's own NUMA node
    eng = capi.Engine(local)
    stream = torch.cuda.Stream()  # a real (non-default) stream: the engine launches on it, the events time it
    torch.cuda.set_stream(stream)
    eng.set_stream(stream.cuda_stream)
    eng.upload_float(tab["offsets"], tab["mean"], tab["isr"], tab["cov_idx"], tab["w2"], tab["log_norm"],
                     mode=capi.MODE_MAX, kernel=capi.KERNEL_AUTO)
    split = eng.float_split
    kname = {capi.KERNEL_DIRECT: "direct_kernel (CUDA cores, reference order)",
             capi.KERNEL_TENSOR: "tc_score_kernel8 (tcgen05: fp16 hi.hi + " +
                                 ("one e4m3 product for both corrections (lo parts x 2^8, undone by scale-input-d)" if split == 2 else "fp16 lo.hi + hi.lo") +
                                 ", fused per-mixture min)"}[eng.float_kernel]

    # this rank's segments: i -> rank i mod N; two alternating batches so consecutive steps read fresh inputs
    seg_ids = [rank + world * i for i in range(2 * segs)]
    x_host = torch.from_numpy(segment_features(ms, seg_ids)).pin_memory()
    xa, xb = x_host[:frames].cuda(), x_host[frames:].cuda()
    # score block [frames][ld]: rows padded to a multiple of 4 floats (ld >= M is the API's layout, include/gmm_b200.h),
    # so every row piece is 16-byte aligned and the kernel's lanes store them with 128-bit stores
    ld = (ms.M + 3) // 4 * 4
    scores = torch.empty((frames, ld), dtype=torch.float32, device="cuda")
    state = {"i": 0}

    def step():
        eng.score_device(frames, xa if state["i"] % 2 == 0 else xb, scores, ld=ld)
        state["i"] += 1

    step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    l0 = eng.launch_count
    sec, w0, w1 = time_device(torch, step, args.steps, args.warmup, dist)
    launches = (eng.launch_count - l0) - 0
    launches_timed = launches * args.steps // (args.steps + args.warmup)
    clocks = sampler.stop(w0, w1) if rank == 0 else None
    sec_per_step = sec / args.steps
    value = world * frames / sec_per_step

    # roofline pass: the same steps again with CUDA events bracketing each launch of the scoring kernel
    eng.kernel_timing(True)
    eng.kernel_time()
    for _ in range(args.steps):
        step()
    kernel_ms, kernel_n = eng.kernel_time()
    eng.kernel_timing(False)
    kernel_sec = kernel_ms / 1e3 / max(1, kernel_n)  # per launch; a step is kernel_n / steps launches
    launches_per_step = max(1, kernel_n // max(1, args.steps))

    # end to end through the host API: pinned features in, host scores out, every step
    e2e_frames = min(frames, args.e2e_segments * SEG_FRAMES)
    out_host = torch.empty((e2e_frames, ld), dtype=torch.float32).pin_memory()
    xh = x_host[:e2e_frames].numpy()
    oh = out_host.numpy()

    def e2e_step():
        eng.score_host(xh, want_best=False, out=oh, ld=ld)

    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        e2e_step()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_sec = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e2e_sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_sec = float(t.item())
    e2e_value = world * e2e_frames * e2e_steps / e2e_sec

    # features in, best paths out: scores stay in HBM and feed the Viterbi consumer
    from rasr_b200.align import AlignmentAutomata, linear_hmm
    al_segs = 2 * segs        # both input batches in one call: groups of segments alternate between two lanes
    al_frames = al_segs * SEG_FRAMES
    rng = np.random.default_rng(77 + rank)
    automata = AlignmentAutomata.from_arc_lists(
        [linear_hmm(rng.integers(0, ms.M, 300), loop=3.0, forward=0.0, skip=30.0) for _ in range(al_segs)])
    al_fo = np.arange(al_segs + 1, dtype=np.uint32) * SEG_FRAMES
    xal = x_host[:al_frames].numpy()
    for _ in range(2):
        eng.score_align_host(al_fo, xal, automata, want_states=False)
    if dist is not None:
        dist.barrier()
    al_steps = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(al_steps):
        eng.score_align_host(al_fo, xal, automata, want_states=False)
    al_sec = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([al_sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        al_sec = float(t.item())
    al_value = world * al_frames * al_steps / al_sec

    # the training pass: features in, nothing out -- alignment and E-step on the device
    eng.create_accumulators(tables.mean_index(ms), ms.n_means)
    for _ in range(2):
        eng.score_align_host(al_fo, xal, automata, accumulate=True, want_paths=False)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(al_steps):
        eng.score_align_host(al_fo, xal, automata, accumulate=True, want_paths=False)
        if dist is not None:  # the pass ends with the combine step: one all-reduce of the f64 estimators
            sharded.all_reduce_accumulators(eng, dist)
    eng.sync()
    tr_sec = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([tr_sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tr_sec = float(t.item())
    tr_value = world * al_frames * al_steps / tr_sec

    # the Baum-Welch training pass: compact scores -> forward-backward posteriors -> weighted E-step, on the device
    for _ in range(2):
        eng.score_baum_welch_host(al_fo, xal, automata, accumulate=True)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(al_steps):
        eng.score_baum_welch_host(al_fo, xal, automata, accumulate=True)
        if dist is not None:
            sharded.all_reduce_accumulators(eng, dist)
    eng.sync()
    bw_sec = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([bw_sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        bw_sec = float(t.item())
    bw_value = world * al_frames * al_steps / bw_sec

    # the same alignment call with the front end on the device: 16 base features per frame cross
    # PCIe, window 9 + 45 x 144 matrix product run at the upload (a throughput case: random matrix, random base features)
    lrng = np.random.default_rng(5 + rank)
    eng.set_lda((lrng.standard_normal((ms.D, 144)) / 12.0).astype(np.float32), 16, 9, 4)
    base_host = torch.from_numpy(lrng.standard_normal((al_frames, 16)).astype(np.float32)).pin_memory().numpy()
    for _ in range(2):
        eng.score_align_host(al_fo, base_host, automata, want_states=False)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(al_steps):
        eng.score_align_host(al_fo, base_host, automata, want_states=False)
    lda_sec = time.perf_counter() - t0
    eng.clear_lda()
    if dist is not None:
        t = torch.tensor([lda_sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        lda_sec = float(t.item())
    lda_value = world * al_frames * al_steps / lda_sec

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec_per_step * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": ("f32 (split operands: fp16 hi.hi + e4m3 corrections, f32 accumulate; <= 1e-4 rel. "
                                           "vs the reference's f32)" if split == 2 else
                                           "f32 (fp16x3 split operands, f32 accumulate)") if
            eng.float_kernel == capi.KERNEL_TENSOR else "f32", "data": "synthetic",
            "config": {"workload": "cfg4-slice: 4501 mixtures x 16 densities x 45 dims, float diagonal-maximum, "
                                   f"{segs} segments x {SEG_FRAMES} frames per GPU per step",
                       "frames_per_step_per_gpu": frames, "segment_sharding": "segment i -> rank i mod N, no collective",
                       "l2": f"scores written per step {frames * ld * 4 / 2**20:.0f} MiB > 126 MiB L2; "
                             "inputs alternate between two batches",
                       "score_block": f"[frames][{ld}] f32, rows padded from M = {ms.M} to a multiple of 4",
                       "e2e_frames_per_step_per_gpu": e2e_frames,
                       "kernel": kname},
            "gpu_launches": launches_timed,
            "gaussian_evals_per_s": value * ms.G,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": e2e_frames * ms.D * 4,
                    "d2h_bytes_per_step": e2e_frames * ld * 4, "frames_per_step_per_gpu": e2e_frames,
                    "d2h_GBps_per_gpu": e2e_value / world * ld * 4 / 1e9,
                    "api": "gmm_score_host (C-ABI, pinned host buffers; the call pipelines kernel launches and "
                           "device-to-host copies in chunks)",
                    "bound": "PCIe: every scored frame returns 4 * ld bytes of scores over the GPU's host link"},
            "e2e_alignment": {"value": al_value, "unit": UNIT, "h2d_bytes_per_step": al_frames * ms.D * 4,
                              "d2h_bytes_per_step": al_frames * 4 + al_segs * 4, "frames_per_step_per_gpu": al_frames,
                              "api": "gmm_score_align_host (C-ABI): host features -> scores of the emissions each "
                                     "segment's automaton uses (compact block, kept in HBM) -> Viterbi best path per "
                                     "segment (300-state loop/forward/skip automata) -> host paths; groups of 296 "
                                     "segments alternate between two streams, tables built by a host thread pool",
                              "note": "not the headline e2e: the consumer of the scores runs on the device, so "
                                      "only 4 bytes per frame return"},
            "e2e_training_pass": {"value": tr_value, "unit": UNIT, "h2d_bytes_per_step": al_frames * ms.D * 4,
                                  "d2h_bytes_per_step": 0, "frames_per_step_per_gpu": al_frames,
                                  "api": "gmm_score_align_accumulate_host (C-ABI): host features -> scores -> Viterbi "
                                         "path -> best density of the aligned mixture -> f64 estimator sums, all in "
                                         "HBM; the estimators are read once per pass (gmm_accumulator_read); at N > 1 "
                                         "every step ends with the all-reduce that combines the ranks' estimators"},
            "e2e_alignment_front_end": {
                "value": lda_value, "unit": UNIT, "h2d_bytes_per_step": al_frames * 16 * 4,
                "d2h_bytes_per_step": al_frames * 4 + al_segs * 4, "frames_per_step_per_gpu": al_frames,
                "api": "gmm_set_lda + gmm_score_align_host: 16 base features per frame in, sliding window (9 frames) and "
                       "45 x 144 matrix product on the device, then as e2e_alignment"},
            "e2e_baum_welch_training_pass": {
                "value": bw_value, "unit": UNIT, "h2d_bytes_per_step": al_frames * ms.D * 4, "d2h_bytes_per_step": al_segs * 8,
                "frames_per_step_per_gpu": al_frames,
                "api": "gmm_score_baum_welch_accumulate_host (C-ABI): host features -> compact scores -> forward-backward "
                       "posteriors per frame and emission -> posterior-weighted E-step, all in HBM; only the segments' "
                       "total scores return"},
            "clocks": clocks}
    if world > 1 and not args.no_extra:
        # every rank takes part; rank 0 reports.  A failure here (e.g. peer access unavailable) must not lose the line.
        try:
            cfg5 = cfg5_mixture_sharded(torch, dist, rank, world, local, stream, steps=max(3, min(args.steps, 10)))
        except Exception as e:  # noqa: BLE001
            cfg5 = {"error": f"{type(e).__name__}: {e}"[:300]}
        line["extra"] = {"cfg5_mixture_sharded": cfg5}
    if rank == 0:
        line["roofline"] = float_roofline(ms, frames / launches_per_step, kernel_sec, peaks, kname, sec_per_step,
                                          ncu_traffic("prof_tc_cfg1_max_32768_r02.txt") if launches_per_step > 1 else None,
                                          sustained=sec >= 0.1, launches_per_step=launches_per_step, split=split or 3)
        if world == 1 and not args.no_cpu:
            os.sched_setaffinity(0, all_cpus)  # the CPU baseline uses every host core again
            cores = len(os.sched_getaffinity(0))
            line["cpu_baseline"] = cpu_reference_run(ms, max(512, 96 * cores), "max", 2)
            legs = cpu_reference_legs(ms, cores)
            if legs:
                line["cpu_baseline"]["other_scorers"] = legs
        if world == 1 and not args.no_extra:
            line["extra"] = extra_cases(torch, eng, peaks, args)
        sys.stdout.flush()
        os.dup2(args.stdout_fd, 1)
        print(json.dumps(line))
        sys.stdout.flush()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()
    return 0


def cfg5_mixture_sharded(torch, dist, rank, world, local, stream, steps=10):
    """BASELINE.json configs[4] at N > 1: a 65536 x 64 x 45 model (4.2 M Gaussians), mixtures sharded N ways, every
    rank scores the same T frames against its shard and ends up with all columns.  Kernel alone / score + NCCL
    all-gather / score with its exchange step inside the call (push over peer memory, no NCCL), T = 1024 and 8192.
    Every rank generates only ITS shard (seed 1001 + rank): scores of different shards are independent, so the model
    is the concatenation of the shards."""
    from rasr_b200 import sharded, synth, tables
    M, K, D = 65536, 64, 45
    mp = sharded.shard_width(M, world)
    m0, m1 = sharded.mixture_range(M, rank, world)
    shard = synth.make_model(m1 - m0, K, D, seed=1001 + rank)
    tab = tables.float_tables(shard)
    res = {"model": f"{M} x {K} x {D} ({M * K} Gaussians), {mp} mixtures ({mp * K} Gaussians) per rank", "n_gpus": world}

    def timed_loop(fn, n):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        dist.barrier()
        ev0.record(stream)
        for _ in range(n):
            fn()
        ev1.record(stream)
        torch.cuda.synchronize()
        tt = torch.tensor([ev0.elapsed_time(ev1) / n], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    sc = sharded.MixtureShardedScorer(tab, rank, world, dist, local, total_mixtures=M)
    sc.bind_stream(stream)
    for T in (1024, 8192):
        x = synth.make_features(shard, T, seed=77)  # the same frames on every rank (seeded by rank 0's shard below)
        dx = torch.from_numpy(x).cuda()
        dist.broadcast(dx, 0)
        for _ in range(2):
            g = sc.score(dx)
        ms_ag = timed_loop(lambda: sc.score(dx), steps)
        ms_k = timed_loop(lambda: sc.engine.score_device(T, dx, sc._local, None, ld=sc.Mp), steps)
        ref = sharded.to_frame_major(sc.score(dx), M).clone()
        case = {"ms_kernel_alone": ms_k, "ms_score_plus_nccl_all_gather": ms_ag,
                "all_gather_bus_GBps": (world - 1) * T * mp * 4 / max(1e-9, (ms_ag - ms_k) * 1e-3) / 1e9,
                "exchange_bytes_sent_per_rank": (world - 1) * T * mp * 4}
        best_ms, best_chunks, same_all = 1e30, 0, True
        for chunks in ((2, 4) if T < 4096 else (4, 8)):  # long calls: more, shorter frame ranges (profiles/sweep_push_n4_r02.txt)
            ps = sharded.PushMixtureShardedScorer(tab, rank, world, dist, local, max_frames=T, chunks=chunks,
                                                  total_mixtures=M)
            ps.engine.set_stream(stream.cuda_stream)
            for _ in range(2):
                ps.score(dx)
            ms_p = timed_loop(lambda: ps.score(dx), steps)
            fm = ps.score(dx)
            torch.cuda.synchronize()
            same = torch.tensor([1.0 if torch.equal(fm[:, :M], ref) and not ps.timed_out() else 0.0], device="cuda")
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
            same_all = same_all and bool(same.item() == 1.0)
            case[f"ms_push_{chunks}_ranges"] = ms_p
            if ms_p < best_ms:
                best_ms, best_chunks = ms_p, chunks
            ps.close()
        # overlap across calls: two block sets in turn, a call's last push under the next call's scoring (one step of the
        # loop = one call in steady state)
        pp = sharded.PushMixtureShardedScorer(tab, rank, world, dist, local, max_frames=T, chunks=best_chunks,
                                              total_mixtures=M, n_blocks=2)
        pp.engine.set_stream(stream.cuda_stream)
        turn = [0]
        pp.begin(dx, 0)

        def pipelined_step():
            k = turn[0]
            pp.begin(dx, 1 - k)
            pp.finish(k)
            turn[0] = 1 - k

        for _ in range(2):
            pipelined_step()
        ms_pp = timed_loop(pipelined_step, steps)
        pp.finish(turn[0])
        torch.cuda.synchronize()
        same = torch.tensor([1.0 if torch.equal(pp.blocks[0][:T, :M], ref) and torch.equal(pp.blocks[1][:T, :M], ref)
                             and not pp.timed_out() else 0.0], device="cuda")
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        case.update({"ms_push_pipelined_two_blocks": ms_pp, "pipelined_bit_identical_to_all_gather": bool(same.item() == 1.0),
                     "speedup_pipelined_over_all_gather": ms_ag / ms_pp})
        pp.close()
        case.update({"ms_push_best": best_ms, "push_ranges": best_chunks, "push_bit_identical_to_all_gather": same_all,
                     "frames_per_s_push": T / (best_ms * 1e-3), "frames_per_s_all_gather": T / (ms_ag * 1e-3),
                     "push_egress_GBps_per_rank": (world - 1) * T * mp * 4 / (best_ms * 1e-3) / 1e9,
                     "nvlink_GBps_per_direction": 900.0,
                     "speedup_over_all_gather": ms_ag / best_ms})
        res[f"T{T}"] = case
        del dx, g, ref
    sc.engine.close()
    return res


def timed_kernel(eng, fn, steps):
    """Device time of the scoring kernel per call, averaged over `steps` calls (events inside the library)."""
    eng.kernel_timing(True)
    eng.kernel_time()
    for _ in range(steps):
        fn()
    ms_total, n = eng.kernel_time()
    eng.kernel_timing(False)
    return ms_total / 1e3 / max(1, steps)  # per call (a call may chunk the frames over several launches)


def extra_cases(torch, eng, peaks, args):
    """The other single-GPU configurations of BASELINE.json, kernel-only, for DESIGN.md's tables (not the headline)."""
    from rasr_b200 import capi, synth, tables
    out = {}
    steps, warmup = max(3, min(args.steps, 10)), 3
    # cfg2: same model through the quantized scorer (bit-exact integer path)
    ms, x = synth.make_config("cfg2", T=8 * SEG_FRAMES)
    qt = tables.quant_tables(ms)
    eng.upload_quant(qt["offsets"], qt["qmean"], qt["cov_idx"], qt["qconst"], qt["isr_scaled"], qt["scaling_squared"])
    T = x.shape[0]
    ld = (ms.M + 3) // 4 * 4
    dx = torch.from_numpy(x).cuda()
    dq = torch.empty((T, ld), dtype=torch.int32, device="cuda")
    db = torch.empty((T, ld), dtype=torch.int32, device="cuda")
    dsc = torch.empty((T, ld), dtype=torch.float32, device="cuda")
    qname = {1: "quant_score_kernel (dp4a, CUDA cores)",
             2: "qtc_score_kernel<16> (tcgen05 kind::i8, constants inside the contraction)"}[eng.quant_kernel]
    sec, _, _ = time_device(torch, lambda: eng.quant_score_device(T, dx, dq, db, ld=ld), steps, warmup)
    ksec = timed_kernel(eng, lambda: eng.quant_score_device(T, dx, dq, db, ld=ld), steps)
    out["cfg2_quantized"] = {"frames_per_s": T * steps / sec, "ms_per_step": 1e3 * sec / steps, "frames": T,
                             "outputs": "q (s32) + arg-min (u32)",
                             "roofline": quant_roofline(ms, T, ksec, peaks, kernel_name=qname,
                                                        traffic=ncu_traffic("prof_quant_cfg2_r02.txt"))}
    # what the RASR plugin asks for: the final f32 score only (one array)
    sec1, _, _ = time_device(torch, lambda: eng.quant_score_device(T, dx, None, None, dsc, ld=ld), steps, warmup)
    ksec1 = timed_kernel(eng, lambda: eng.quant_score_device(T, dx, None, None, dsc, ld=ld), steps)
    out["cfg2_quantized_score_only"] = {"frames_per_s": T * steps / sec1, "ms_per_step": 1e3 * sec1 / steps, "frames": T,
                                        "outputs": "final f32 score",
                                        "roofline": quant_roofline(ms, T, ksec1, peaks, outputs=1, kernel_name=qname)}
    del dq, db, dsc
    # cfg3: 12001 x 64, 10k-frame batches, max and sum
    ms, x = synth.make_config("cfg3", T=10000)
    tab = tables.float_tables(ms)
    T = x.shape[0]
    ld = (ms.M + 3) // 4 * 4
    dx = torch.from_numpy(x).cuda()
    ds = torch.empty((T, ld), dtype=torch.float32, device="cuda")
    for mode, name in ((capi.MODE_MAX, "cfg3_max"), (capi.MODE_SUM, "cfg3_sum")):
        eng.upload_float(tab["offsets"], tab["mean"], tab["isr"], tab["cov_idx"], tab["w2"], tab["log_norm"],
                         mode=mode, kernel=capi.KERNEL_AUTO)
        st = steps if eng.float_kernel == capi.KERNEL_TENSOR else 2
        sec, _, _ = time_device(torch, lambda: eng.score_device(T, dx, ds, ld=ld), st, 1 if st == 2 else warmup)
        ksec = timed_kernel(eng, lambda: eng.score_device(T, dx, ds, ld=ld), st)
        prof = {"cfg3_max": "prof_tc_cfg3_max_10000_r02.txt", "cfg3_sum": "prof_tc_cfg3_sum_10000_r02.txt"}[name]
        kn = {capi.KERNEL_DIRECT: "direct_kernel", capi.KERNEL_TENSOR:
              ("tc_score_kernel8" if mode == capi.MODE_MAX else "tc_score_kernel16") + f" ({eng.float_split} MMA products)"}
        out[name] = {"frames_per_s": T * st / sec, "ms_per_step": 1e3 * sec / st, "frames": T,
                     "kernel": eng.float_kernel,
                     "roofline": float_roofline(ms, T, ksec, peaks, kn[eng.float_kernel], sec / st, ncu_traffic(prof),
                                                split=eng.float_split or 3)}
    if not args.no_cpu:
        cores = len(os.sched_getaffinity(0))
        legs = cpu_reference_legs(ms, cores, scorers=("max", "sum"))
        if legs:
            out["cfg3_cpu_reference"] = legs
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--segments", type=int, default=512, help="1000-frame segments per GPU per step")
    ap.add_argument("--e2e-segments", type=int, default=32, help="segments per host-API call of the e2e leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the cfg2 / cfg3 side measurements")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)
    # stdout carries the one JSON line only: whatever native libraries print there meanwhile (NCCL's version banner at
    # communicator creation) is sent to stderr; the descriptor is restored just before the line is printed
    sys.stdout.flush()
    args.stdout_fd = os.dup(1)
    os.dup2(2, 1)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
