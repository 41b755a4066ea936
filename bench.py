#!/usr/bin/env python
"""Throughput benchmark of the fused spectral front end (BASELINE.json metric: audio-seconds per second).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload logfiltspecflux|superflux|multires_beat|
                    multires_onset] [--tracks 1024] [--seconds 30] [--impl ours|reference]

One *step* = the whole batch (default: BASELINE.json configs[1], LogFiltSpecFlux features -- the (T, 81)
positive difference matrix and the (T,) flux -- for 1024 synthetic 30 s 44.1 kHz int16 tracks per GPU)
through the hot path once.  Prints ONE JSON line (rank 0):

  value     audio-s/s, whole job, PCM already resident in HBM (one fused kernel launch per recipe and step)
  e2e       the same through the host API (BatchFrontEnd.run_host -> mmb_run_host): pinned host PCM in, pinned
            host features out, H2D / kernel / D2H inside the timed region
  roofline  of the fused kernel, per launch, from CUDA events on the launching stream: algorithmic HBM bytes
            (hop*sizeof(sample) in + feature columns out per frame) / time vs MEASURED_PEAKS.json hbm_gbs;
            `fp32` carries the FP32-pipe roof (canonical flop count / measured FFMA peak), which is the one
            that binds for this path
  cpu_baseline  the numpy oracle (port of the reference's own numpy/scipy path) on the host cores, bounded
            sample, multiprocessing over tracks like the reference's process_batch

`--impl reference` times only that CPU path (rank 0; other ranks exit) and prints the same line shape.
Multi-GPU: tracks are sharded over ranks with no collective (weak scaling: 1024 tracks per GPU).
"""
import argparse
import json
import math
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SR = 44100


# ----------------------------------------------------------------------------------------------- workloads
def workload_recipes(name):
    from madmom_b200 import recipe as R
    if name == "logfiltspecflux":
        return [R.LOGFILTSPECFLUX], ("diff", "odf")
    if name == "superflux":
        return [R.SUPERFLUX], ("diff",)
    if name == "multires_beat":
        return list(R.RNN_BEAT), ("stack",)
    if name == "multires_onset":
        return list(R.RNN_ONSET), ("stack",)
    raise SystemExit("unknown workload %r" % name)


def flops_per_frame(rec, bands, nnz):
    """Canonical count of BASELINE.md section 4: 2.5 N log2 N + N + 4 K + 2 nnz + 3 B [+ 2 B] + 2 B."""
    N = rec.frame_size
    f = 2.5 * N * math.log2(N) + N + 4 * (N // 2) + 2 * nnz + 3 * bands + 2 * bands
    if rec.diff_max_bins and rec.diff_max_bins > 1:
        f += 2 * bands
    return f


# ----------------------------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML every 100 ms while running."""
    BITS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
            0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, index):
        threading.Thread.__init__(self, daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        while not self._stop_evt.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.BITS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.1)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(2.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ----------------------------------------------------------------------------------------------- CPU arm
def _cpu_init():
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:
        pass


def _cpu_task(args):
    """One track through the numpy oracle; returns processing seconds (generation excluded)."""
    workload, seed, seconds = args
    from madmom_b200.synth import make_track
    from oracle import frontend as O
    x = make_track(seed, seconds)
    t0 = time.perf_counter()
    if workload == "logfiltspecflux":
        st = O.run_recipe(x, O.LOGFILTSPECFLUX, stages=True)
        _ = st["diff"], st["odf"]
    elif workload == "superflux":
        O.run_recipe(x, O.SUPERFLUX)
    else:
        mul, ratio = (1, 0.5) if workload == "multires_beat" else (5, 0.25)
        cols = []
        for N in (1024, 2048, 4096):
            st = O.run_recipe(x, dict(frame_size=N, hop_size=441., num_bands=6, norm_filters=True, mul=mul, add=1,
                                      diff_ratio=ratio, positive_diffs=True), stages=True)
            cols += [st["log"], st["diff"]]
        np.hstack(cols)
    return os.getpid(), time.perf_counter() - t0


def cpu_reference_throughput(workload, seconds, budget_s=20.0, cores=None):
    """audio-s/s of the CPU path using a multiprocessing.Pool over tracks (the reference's own process_batch
    strategy).  Bounded: about `budget_s` seconds of wall time."""
    import multiprocessing as mp
    cores = cores or len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    _cpu_init()
    t_one = _cpu_task((workload, 999, min(seconds, 10.0)))[1] * (seconds / min(seconds, 10.0))   # warm imports
    per_core = min(64, max(1, int(round(budget_s / max(t_one, 1e-3)))))
    n_tracks = cores * per_core
    ctx = mp.get_context("fork")
    with ctx.Pool(cores, initializer=_cpu_init) as pool:
        pool.map(_cpu_task, [(workload, 1, 0.5)] * cores)                                      # spin up workers
        res = pool.map(_cpu_task, [(workload, 1000 + (i % 8), seconds) for i in range(n_tracks)], chunksize=1)
    by_worker = {}
    for pid, dt in res:
        by_worker[pid] = by_worker.get(pid, 0.0) + dt
    wall = max(by_worker.values())            # processing time of the busiest worker (track synthesis excluded)
    busy = sum(by_worker.values())
    audio = n_tracks * seconds
    return {"value": audio / wall, "unit": "audio-s/s", "cores": cores, "kind": "port",
            "sample": "%d x %.0f s seeded synthetic tracks (8 distinct seeds), numpy oracle "
                      "(scipy.fftpack float64 FFT), multiprocessing.Pool(%d) over tracks, BLAS threads 1; "
                      "%.1f s on the busiest worker; single-core %.1f audio-s/s"
                      % (n_tracks, seconds, cores, wall, audio / max(busy, 1e-9)),
            "single_core_value": audio / max(busy, 1e-9)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    base = cpu_reference_throughput(args.workload, args.seconds, budget_s=max(5.0, 12.0 * max(args.steps, 1) / 5.0))
    line = {"impl": "reference", "metric": "audio-seconds per second (x realtime), spectral front end",
            "value": base["value"], "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "tracks": args.tracks, "seconds": args.seconds,
                       "note": "CPU path of the reference (numpy/scipy port in oracle/), all host cores"},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from madmom_b200 import _lib
    from madmom_b200.batch import BatchFrontEnd
    from madmom_b200.synth import make_batch_torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: madmom_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_bound = False
    if world > 1:
        from madmom_b200 import parallel
        numa_bound = parallel.bind_host_thread_to_gpu(local) if os.environ.get("MMB_NO_NUMA_BIND") is None else False
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    recipes, outputs = workload_recipes(args.workload)
    fe = BatchFrontEnd(recipes, np.int16, device=local)
    n_samples = int(round(args.seconds * SR))
    # rank r owns tracks [r * tracks, (r + 1) * tracks): independent units, no collective
    pcm = make_batch_torch(args.tracks, args.seconds, device=dev, seed=1000 + rank)
    lengths = [n_samples] * args.tracks
    units, T = fe.units(lengths)
    rows = int(sum(T))
    frames_per_step = rows
    audio_per_step = rows / recipes[0].fps

    # ---------------- device-resident arm
    out = fe.run_device(pcm, lengths, outputs, units=units, rows=rows)
    torch.cuda.synchronize()
    launches_per_step = len(recipes)
    for _ in range(args.warmup):
        fe.run_device(pcm, lengths, outputs, out=out, units=units, rows=rows)
    sampler = ClockSampler(local)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    l0 = _lib.launch_count
    barrier()
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for a, b in ev:
        a.record()
        fe.run_device(pcm, lengths, outputs, out=out, units=units, rows=rows)
        b.record()
    t1.record()
    barrier()
    ms_total = max_over_ranks(t0.elapsed_time(t1))
    step_ms = [a.elapsed_time(b) for a, b in ev]
    gpu_launches = _lib.launch_count - l0
    ms_per_step = ms_total / args.steps
    value = world * audio_per_step / (ms_per_step * 1e-3)

    # ---------------- end-to-end arm (host buffers through the C ABI host path)
    e2e = None
    if not args.no_e2e:
        pcm_host = torch.empty(pcm.shape, dtype=pcm.dtype, pin_memory=True)
        pcm_host.copy_(pcm)
        torch.cuda.synchronize()
        e2e_steps = max(2, min(args.steps, 5))
        host_out = fe.run_host(pcm_host, lengths, outputs)
        barrier()
        w0 = time.perf_counter()
        for _ in range(e2e_steps):
            fe.run_host(pcm_host, lengths, outputs, out=host_out)
        torch.cuda.synchronize()
        e2e_s = max_over_ranks(time.perf_counter() - w0) / e2e_steps
        e2e = {"value": world * audio_per_step / e2e_s, "unit": "audio-s/s",
               "h2d_bytes_per_step": pcm_host.numel() * pcm_host.element_size() * len(recipes),
               "d2h_bytes_per_step": sum(t.numel() * t.element_size() for t in host_out.values()),
               "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
               "api": "BatchFrontEnd.run_host -> mmb_run_host (pinned host in/out)"}
    clocks = sampler.finish()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the fused kernel (per launch, live CUDA events)
    sample_bytes = 2
    cols = sum(t.shape[1] if t.dim() == 2 else 1 for t in out.values())
    bytes_per_frame = recipes[0].hop_size * sample_bytes * len(recipes) + cols * 4
    fpf = 0.0
    for r, plan in zip(recipes, fe.plans):
        s, l, w = r.build_filterbank().band_table()
        fpf += flops_per_frame(r, plan.num_bands, int(l.sum()))
    launch_ms = float(np.mean(step_ms)) / launches_per_step
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
        hbm_peak, peak_src = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        hbm_peak, peak_src = 6650.0, "fallback"
    fp32_peak = _lib.measure_fp32_peak(local)
    ach_gbs = frames_per_step * bytes_per_frame / (float(np.mean(step_ms)) * 1e-3) / 1e9
    ach_tf = frames_per_step * fpf / (float(np.mean(step_ms)) * 1e-3) / 1e12
    traffic = counts = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            with open(tp) as f:
                t = json.load(f).get(args.workload)
            # measured DRAM bytes per frame (ncu --set full of the same command) x frames of one launch
            traffic = float(t["dram_bytes_per_frame"]) * frames_per_step / len(recipes) if t else None
            counts = t if t and "warp_instructions_per_frame" in t else None
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
                "traffic": traffic, "peak_source": peak_src, "kernel": "mmb::frontend_kernel",
                "launch_ms": launch_ms, "algorithmic_bytes_per_frame": bytes_per_frame,
                "frames_per_launch": frames_per_step,
                "fp32": {"binding": True, "achieved": ach_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": ach_tf / fp32_peak if fp32_peak else None, "flops_per_frame": fpf,
                         "peak_source": "measured (mmb_measure_fp32_peak, dependent FFMA chains, best of 5)"}}

    if counts and clocks.get("sm_mhz"):
        # the ceilings that actually bind (DESIGN.md "Which roof binds"): 4 warp-instructions and 1 shared-memory
        # wavefront per clock and SM (measured, mmb_rate_probe); counts per frame from the committed ncu capture,
        # time and clock from this run
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        clk_per_frame = launch_ms * 1e-3 * clocks["sm_mhz"] * 1e6 * sms / (frames_per_step / len(recipes))
        roofline["sm"] = {
            "sm_clocks_per_frame": clk_per_frame,
            "issue_slots": {"per_frame": counts["warp_instructions_per_frame"], "peak_per_clk_sm": 4.0,
                            "frac": counts["warp_instructions_per_frame"] / 4.0 / clk_per_frame},
            "shared_pipe": {"per_frame": counts["shared_wavefronts_per_frame"], "peak_per_clk_sm": 1.0,
                            "frac": counts["shared_wavefronts_per_frame"] / clk_per_frame},
            "counts_source": counts.get("counts_source")}

    cpu = None
    if world == 1 and not args.no_cpu:
        cpu = cpu_reference_throughput(args.workload, args.seconds, budget_s=15.0)

    line = {"metric": "audio-seconds per second (x realtime), spectral front end", "value": value,
            "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "tracks_per_gpu": args.tracks, "seconds": args.seconds,
                       "sample_rate": SR, "pcm": "int16", "frames_per_step_per_gpu": frames_per_step,
                       "outputs": {k: list(v.shape) for k, v in out.items()},
                       "l2": "inputs (%.2f GB PCM per step) larger than L2" % (pcm.numel() * 2 / 1e9),
                       "parallelism": "tracks sharded over %d GPU(s), no collective" % world,
                       "host_numa_bound": numa_bound},
            "e2e": e2e,
            "gpu_launches": gpu_launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="logfiltspecflux")
    ap.add_argument("--tracks", type=int, default=1024, help="tracks per GPU")
    ap.add_argument("--seconds", type=float, default=30.0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host end-to-end leg (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
