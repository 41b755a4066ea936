#!/usr/bin/env python
"""Throughput benchmark of the fused spectral front end (BASELINE.json metric: audio-seconds per second).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload logfiltspecflux|superflux|multires_beat|
                    multires_onset|complexflux] [--tracks 1024] [--seconds 30] [--impl ours|reference]

One *step* = the whole batch (default: BASELINE.json configs[1], LogFiltSpecFlux features -- the (T, 81)
positive difference matrix and the (T,) flux -- for 1024 synthetic 30 s 44.1 kHz int16 tracks per GPU)
through the hot path once.  Prints ONE JSON line (rank 0):

  value        audio-s/s, whole job, PCM already resident in HBM (one fused kernel launch per step)
  e2e          the same through the host API (BatchFrontEnd.run_host -> mmb_run_host): pinned host PCM in, pinned
               host features out, H2D / kernel / D2H inside the timed region
  e2e_program  host PCM in -> flux -> peak picking on the device -> onset frame indices out (the LogFiltSpecFlux
               program; only the peaks travel back)
  roofline     of the fused kernel per launch from CUDA events on the launching stream.  The binding roof is the
               FP32 pipe: canonical flop count / measured FFMA peak at the top level, the HBM figures
               (algorithmic bytes / MEASURED_PEAKS.json hbm_gbs, ncu traffic) nested under `hbm`, the measured
               FFT-only ceiling of the same warp FFT under `fft_only` (+ `fft_only_frac`)
  configs      (N = 1) short timings of the other BASELINE configurations with both fractions
  long_signal  (N > 1) BASELINE config 4 on a shortened signal: NCCL all-gather vs the fused peer-store / multicast
               kernels, bit-identity of the three and of the unsharded launch
  host_link    aggregate pinned-copy GB/s of all ranks copying at once: H2D only, D2H only, both; the e2e blocks
               carry their own H2D rate and its fraction of this link (`h2d_gbs`, `link_frac`)
  cpu_baseline the reference's own classes (copy of the reference under baseline/_ref, `kind: reference`) or the
               numpy port in oracle/ (`kind: port`) on the host cores, bounded sample, multiprocessing over tracks

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
REF_COPY = os.path.join(ROOT, "baseline", "_ref")


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
    if name == "complexflux":
        return [R.COMPLEXFLUX], ("odf",)
    raise SystemExit("unknown workload %r" % name)


def flops_per_frame(rec, bands, nnz, complexflux=False):
    """Canonical count of SURVEY.md 8(d): 2.5 N log2 N + N + 4 K + 2 nnz + 3 B [+ 2 B] + 2 B; ComplexFlux adds
    about 10 K + nnz for the phase differences, the temporal maximum and the band minima."""
    N = rec.frame_size
    f = 2.5 * N * math.log2(N) + N + 4 * (N // 2) + 2 * nnz + 3 * bands + 2 * bands
    if rec.diff_max_bins and rec.diff_max_bins > 1:
        f += 2 * bands
    if complexflux:
        f += 10 * (N // 2) + nnz
    return f


def config_dict(args, **extra):
    """The keys both arms print, so that the driver can tell they ran the same workload."""
    d = {"workload": args.workload, "tracks_per_gpu": args.tracks, "seconds": args.seconds, "pcm": "int16",
         "sample_rate": SR}
    d.update(extra)
    return d


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
def reference_copy_available():
    return os.path.isdir(os.path.join(REF_COPY, "madmom", "audio"))


def _cpu_init():
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:
        pass


_REF_PROCS = {}


def _reference_processor(workload):
    """The reference's own processor chain for a workload, built once per worker process."""
    proc = _REF_PROCS.get(workload)
    if proc is not None:
        return proc
    os.environ["MADMOM_REFERENCE_ROOT"] = REF_COPY
    from oracle.ref_loader import load_reference
    ref = load_reference()
    ON, FI, SP, S, ST, P = ref.onsets, ref.filters, ref.spectrogram, ref.signal, ref.stft, ref.processors
    if workload in ("logfiltspecflux", "superflux", "complexflux"):
        # bin/LogFiltSpecFlux:60-69, bin/SuperFlux:59-69, bin/ComplexFlux:60-74 -> SpectralOnsetProcessor
        # (madmom/features/onsets.py:651-684)
        kw = dict(sample_rate=SR, filterbank=FI.LogarithmicFilterbank, fmin=30, fmax=17000, norm_filters=False,
                  log=np.log10, mul=1, add=1, diff_ratio=0.5, positive_diffs=True)
        if workload == "logfiltspecflux":
            proc = ON.SpectralOnsetProcessor(onset_method="spectral_flux", fps=100, num_bands=12, **kw)
        elif workload == "superflux":
            proc = ON.SpectralOnsetProcessor(onset_method="superflux", fps=100, num_bands=24, diff_max_bins=3, **kw)
        else:
            proc = ON.SpectralOnsetProcessor(onset_method="complex_flux", fps=200, num_bands=24, diff_max_bins=3, **kw)
    else:
        # RNNBeatProcessor / RNNOnsetProcessor pre-processing (features/beats.py:93-106, features/onsets.py:778-791)
        mul, ratio = (1, 0.5) if workload == "multires_beat" else (5, 0.25)
        multi = P.ParallelProcessor([])
        for frame_size in [1024, 2048, 4096]:
            frames = S.FramedSignalProcessor(frame_size=frame_size, fps=100)
            stft = ST.ShortTimeFourierTransformProcessor()
            filt = SP.FilteredSpectrogramProcessor(num_bands=6, fmin=30, fmax=17000, norm_filters=True)
            spec = SP.LogarithmicSpectrogramProcessor(mul=mul, add=1)
            diff = SP.SpectrogramDifferenceProcessor(diff_ratio=ratio, positive_diffs=True, stack_diffs=np.hstack)
            multi.append(P.SequentialProcessor((frames, stft, filt, spec, diff)))
        proc = P.SequentialProcessor((S.SignalProcessor(num_channels=1, sample_rate=SR), multi, np.hstack))
    _REF_PROCS[workload] = proc
    return proc


def _cpu_task(args):
    """One track through the CPU path; returns (pid, processing seconds) -- track synthesis excluded."""
    kind, workload, seed, seconds = args
    from madmom_b200.synth import make_track
    x = make_track(seed, seconds)
    if kind == "reference":
        proc = _reference_processor(workload)
        t0 = time.perf_counter()
        np.asarray(proc(x))
        return os.getpid(), time.perf_counter() - t0
    from oracle import frontend as O
    t0 = time.perf_counter()
    if workload == "logfiltspecflux":
        st = O.run_recipe(x, O.LOGFILTSPECFLUX, stages=True)
        _ = st["diff"], st["odf"]
    elif workload == "superflux":
        O.run_recipe(x, O.SUPERFLUX)
    elif workload == "complexflux":
        st = O.run_recipe(x, dict(frame_size=2048, hop_size=220.5, num_bands=24, norm_filters=False, mul=1, add=1,
                                  diff_ratio=0.5, diff_max_bins=3, positive_diffs=True, circular_shift=True),
                          stages=True)
        O.complex_flux(st["log"], st["stft"], st["filterbank"], st["diff_frames"])
    else:
        mul, ratio = (1, 0.5) if workload == "multires_beat" else (5, 0.25)
        cols = []
        for N in (1024, 2048, 4096):
            st = O.run_recipe(x, dict(frame_size=N, hop_size=441., num_bands=6, norm_filters=True, mul=mul, add=1,
                                      diff_ratio=ratio, positive_diffs=True), stages=True)
            cols += [st["log"], st["diff"]]
        np.hstack(cols)
    return os.getpid(), time.perf_counter() - t0


def cpu_throughput(workload, seconds, kind, budget_s=20.0, cores=None):
    """audio-s/s of the CPU path with a multiprocessing.Pool over tracks (the reference's own process_batch
    strategy, madmom/processors.py:637-713).  Bounded: about `budget_s` seconds of wall time."""
    import multiprocessing as mp
    cores = cores or (len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    _cpu_init()
    if kind == "reference":
        os.environ["MADMOM_REFERENCE_ROOT"] = REF_COPY
    short = min(seconds, 10.0)
    _cpu_task((kind, workload, 999, short))                                     # imports, processor construction
    t_one = _cpu_task((kind, workload, 999, short))[1] * (seconds / short)
    per_core = min(64, max(1, int(round(budget_s / max(t_one, 1e-3)))))
    n_tracks = cores * per_core
    ctx = mp.get_context("fork")
    with ctx.Pool(cores, initializer=_cpu_init) as pool:
        pool.map(_cpu_task, [(kind, workload, 1, 0.5)] * cores)                 # spin up workers
        w0 = time.perf_counter()
        res = pool.map(_cpu_task, [(kind, workload, 1000 + (i % 8), seconds) for i in range(n_tracks)], chunksize=1)
        pool_wall = time.perf_counter() - w0
    by_worker = {}
    for pid, dt in res:
        by_worker[pid] = by_worker.get(pid, 0.0) + dt
    wall = max(by_worker.values())            # processing time of the busiest worker (track synthesis excluded)
    busy = sum(by_worker.values())
    audio = n_tracks * seconds
    what = ("the reference's own classes (copy of CPJKU/madmom under baseline/_ref, scipy.fftpack float64 FFT)"
            if kind == "reference" else "numpy port of the reference path (oracle/frontend.py, scipy.fftpack float64 FFT)")
    return {"value": audio / wall, "unit": "audio-s/s", "cores": cores, "kind": kind,
            "sample": "%d x %.0f s seeded synthetic tracks (8 distinct seeds) through %s, multiprocessing.Pool(%d) "
                      "over tracks, BLAS threads 1; %.1f s on the busiest worker (%.1f s pool wall incl. track "
                      "synthesis); single-core %.1f audio-s/s"
                      % (n_tracks, seconds, what, cores, wall, pool_wall, audio / max(busy, 1e-9)),
            "single_core_value": audio / max(busy, 1e-9)}


def cpu_baseline(workload, seconds, budget_s):
    """`kind: reference` when the copy of the reference is present, with the port's number beside it."""
    if reference_copy_available():
        base = cpu_throughput(workload, seconds, "reference", budget_s)
        try:
            port = cpu_throughput(workload, seconds, "port", max(4.0, budget_s / 3.0))
            base["port_value"], base["port_single_core_value"] = port["value"], port["single_core_value"]
        except Exception as exc:                                                # the port is only a side note
            base["port_value"] = None
            base["port_error"] = str(exc)[:200]
        return base
    base = cpu_throughput(workload, seconds, "port", budget_s)
    base["note"] = "baseline/_ref (copy of the reference's Python sources) not present: timed the oracle port"
    return base


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    base = cpu_baseline(args.workload, args.seconds, budget_s=max(5.0, 12.0 * max(args.steps, 1) / 5.0))
    line = {"impl": "reference", "metric": "audio-seconds per second (x realtime), spectral front end",
            "value": base["value"], "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, note="CPU path of the reference on all host cores; each step a bounded "
                                             "sample of the workload (see cpu_baseline.sample)"),
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------- GPU arm
def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def _traffic(workload):
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(tp) as f:
            return json.load(f).get(workload)
    except Exception:
        return None


def roofline_of(recipes, plans, frames, ms, launches, fp32_peak, hbm_peak, peak_src, out_cols, workload,
                fft_only_fps=None, clocks=None, sms=None):
    """Both roofs of one step: FP32 (binding) at the top level, HBM nested."""
    cf = workload == "complexflux"
    fpf = 0.0
    for r, plan in zip(recipes, plans):
        s, l, w = r.build_filterbank().band_table()
        fpf += flops_per_frame(r, plan.num_bands, int(l.sum()), cf)
    bytes_per_frame = recipes[0].hop_size * 2 + out_cols * 4            # PCM counted once
    sec = ms * 1e-3
    ach_tf = frames * fpf / sec / 1e12
    ach_gbs = frames * bytes_per_frame / sec / 1e9
    t = _traffic(workload)
    traffic = float(t["dram_bytes_per_frame"]) * frames if t and "dram_bytes_per_frame" in t else None
    roof = {"bound": "fp32", "achieved": ach_tf, "peak": fp32_peak, "unit": "TFLOP/s",
            "frac": ach_tf / fp32_peak if fp32_peak else None, "flops_per_frame": fpf,
            "traffic": traffic, "kernel": "mmb::frontend_kernel", "launch_ms": ms / max(launches, 1),
            "launches_per_step": launches, "frames_per_step": frames,
            "peak_source": "measured (mmb_measure_fp32_peak, dependent FFMA chains, best of 5)",
            "hbm": {"achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
                    "algorithmic_bytes_per_frame": bytes_per_frame, "traffic": traffic, "peak_source": peak_src}}
    if fft_only_fps:
        roof["fft_only"] = {"frames_per_s": fft_only_fps,
                            "tflops": fft_only_fps * (2.5 * 2048 * 11 + 4 * 1024) / 1e12,
                            "frac_of_fp32_peak": fft_only_fps * (2.5 * 2048 * 11 + 4 * 1024) / 1e12 / fp32_peak,
                            "what": "mmb_fft_probe: the kernel's 2048-point warp FFT + split + magnitudes alone, "
                                    "16 warps/SM, no staging / window / filterbank / stores"}
        roof["fft_only_frac"] = (frames / sec) / fft_only_fps
    if t and "warp_instructions_per_frame" in t and clocks and clocks.get("sm_mhz") and sms:
        # the ceilings that actually bind (DESIGN.md "Which roof binds"): 4 warp-instructions and 1 shared-memory
        # wavefront per clock and SM (measured, mmb_rate_probe); counts per frame from the committed ncu capture
        clk = sec * clocks["sm_mhz"] * 1e6 * sms / frames
        roof["sm"] = {"sm_clocks_per_frame": clk,
                      "issue_slots": {"per_frame": t["warp_instructions_per_frame"], "peak_per_clk_sm": 4.0,
                                      "frac": t["warp_instructions_per_frame"] / 4.0 / clk},
                      "shared_pipe": {"per_frame": t["shared_wavefronts_per_frame"], "peak_per_clk_sm": 1.0,
                                      "frac": t["shared_wavefronts_per_frame"] / clk},
                      "counts_source": t.get("counts_source")}
    return roof


def run_ours(args):
    import torch
    import torch.distributed as dist
    from madmom_b200 import _lib
    from madmom_b200.batch import BatchFrontEnd, complex_flux_batch
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

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    n_samples = int(round(args.seconds * SR))
    # rank r owns tracks [r * tracks, (r + 1) * tracks): independent units, no collective
    pcm = make_batch_torch(args.tracks, args.seconds, device=dev, seed=1000 + rank)
    lengths = [n_samples] * args.tracks

    def time_workload(name, steps, warmup):
        """Device-resident timing of one workload: (ms per step max over ranks, per-step event ms, launches, frames,
        outputs, front end)."""
        recipes, outputs = workload_recipes(name)
        fe = BatchFrontEnd(recipes, np.int16, device=local)
        units, T = fe.units(lengths)
        rows = int(sum(T))
        if name == "complexflux":
            step = lambda out: {"odf": complex_flux_batch(fe, pcm, lengths)}
            out = step(None)
        else:
            out = fe.run_device(pcm, lengths, outputs, units=units, rows=rows)
            step = lambda out: fe.run_device(pcm, lengths, outputs, out=out, units=units, rows=rows)
        torch.cuda.synchronize()
        for _ in range(warmup):
            step(out)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        l0 = _lib.launch_count
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for a, b in ev:
            a.record()
            step(out)
            b.record()
        t1.record()
        barrier()
        ms = max_over_ranks(t0.elapsed_time(t1)) / steps
        return ms, [a.elapsed_time(b) for a, b in ev], _lib.launch_count - l0, rows, out, fe

    # ---------------- device-resident arm
    sampler = ClockSampler(local)
    sampler.start()
    ms_per_step, step_ms, gpu_launches, rows, out, fe = time_workload(args.workload, args.steps, args.warmup)
    recipes = fe.recipes
    audio_per_step = rows / recipes[0].fps
    value = world * audio_per_step / (ms_per_step * 1e-3)
    outputs = tuple(out.keys())

    # ---------------- end-to-end arms (host buffers through the C ABI host path)
    e2e = e2e_program = None
    pcm_host = None
    if not args.no_e2e and args.workload != "complexflux":
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
        if args.workload in ("logfiltspecflux", "superflux"):
            # the whole program: only the onsets come back (bin/LogFiltSpecFlux:68-70, bin/SuperFlux:67-69)
            thr = 1.6 if args.workload == "logfiltspecflux" else 1.1
            kw = dict(threshold=thr, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0.)
            peaks, counts, _ = fe.onsets_host(pcm_host, lengths, **kw)
            barrier()
            w0 = time.perf_counter()
            for _ in range(e2e_steps):
                peaks, counts, _ = fe.onsets_host(pcm_host, lengths, **kw)
            torch.cuda.synchronize()
            p_s = max_over_ranks(time.perf_counter() - w0) / e2e_steps
            e2e_program = {"value": world * audio_per_step / p_s, "unit": "audio-s/s", "ms_per_step": p_s * 1e3,
                           "h2d_bytes_per_step": pcm_host.numel() * pcm_host.element_size(),
                           "d2h_bytes_per_step": int(peaks.numel() * 4 + counts.numel() * 4),
                           "onsets_per_step": int(peaks.numel()), "steps": e2e_steps,
                           "api": "BatchFrontEnd.onsets_host: pinned PCM -> H2D -> fused kernel (flux only) -> "
                                  "mmb_peak_picking on the device -> peak frame indices to the host"}
    clocks = sampler.finish()

    # ---------------- N > 1: host link probe and BASELINE config 4 (long signal, halo + all-gather)
    host_link = long_signal = None
    if not args.no_e2e:
        host_link = probe_host_link(torch, dev, barrier, max_over_ranks, sum_over_ranks)
    if world > 1:
        try:
            long_signal = run_long_signal(torch, dist, dev, local, world, rank, barrier, max_over_ranks)
        except Exception as exc:                                                 # never lose the headline line
            long_signal = {"error": "%s: %s" % (type(exc).__name__, str(exc)[:300])}

    # ---------------- N = 1: the other BASELINE configurations, short timings
    fp32_peak = _lib.measure_fp32_peak(local)
    hbm_peak, peak_src = _peaks()
    sms = torch.cuda.get_device_properties(local).multi_processor_count
    configs = None
    if world == 1 and not args.no_configs:
        configs = {}
        del out
        for name in ("superflux", "multires_beat", "multires_onset", "complexflux"):
            if name == args.workload:
                continue
            try:
                ms, sm, nl, fr, o, f2 = time_workload(name, 4, 3)
                cols = sum(t.shape[1] if t.dim() == 2 else 1 for t in o.values())
                roof = roofline_of(f2.recipes, f2.plans, fr, float(np.mean(sm)), nl // 4, fp32_peak, hbm_peak,
                                   peak_src, cols, name)
                configs[name] = {"ms_per_step": ms, "value": fr / f2.recipes[0].fps / (ms * 1e-3),
                                 "gpu_launches_per_step": nl // 4, "frames_per_step": fr,
                                 "fp32_frac": roof["frac"], "hbm_frac": roof["hbm"]["frac"],
                                 "dram_traffic_bytes_per_step": roof["traffic"],
                                 "flops_per_frame": roof["flops_per_frame"],
                                 "algorithmic_bytes_per_frame": roof["hbm"]["algorithmic_bytes_per_frame"],
                                 "outputs": {k: list(v.shape) for k, v in o.items()}}
                del o, f2
                torch.cuda.empty_cache()
            except Exception as exc:
                configs[name] = {"error": "%s: %s" % (type(exc).__name__, str(exc)[:300])}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the fused kernel (per launch, live CUDA events)
    fft_fps = None
    try:
        fft_fps = _lib.fft_only_frames_per_s(local)
    except Exception:
        fft_fps = None
    _, outs_spec = workload_recipes(args.workload)
    cols = {"diff": fe.plans[0].num_bands, "odf": 1, "stack": fe.stack_width}
    out_cols = sum(cols.get(n, 0) for n in outs_spec)
    launches_per_step = max(gpu_launches // max(args.steps, 1), 1)
    roofline = roofline_of(recipes, fe.plans, rows, float(np.mean(step_ms)), launches_per_step, fp32_peak, hbm_peak,
                           peak_src, out_cols, args.workload, fft_fps if args.workload != "complexflux" else None,
                           clocks, sms)

    cpu = None
    if world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args.workload, args.seconds, budget_s=15.0)

    line = {"metric": "audio-seconds per second (x realtime), spectral front end", "value": value,
            "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_dict(args, frames_per_step_per_gpu=rows, outputs=list(outs_spec),
                                  l2="inputs (%.2f GB PCM per step) larger than L2" % (pcm.numel() * 2 / 1e9),
                                  parallelism="tracks sharded over %d GPU(s), no collective" % world,
                                  host_numa_bound=numa_bound),
            "e2e": e2e, "e2e_program": e2e_program,
            "gpu_launches": gpu_launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu}
    if configs is not None:
        line["configs"] = configs
    if host_link is not None:
        line["host_link"] = host_link
        for key in ("e2e", "e2e_program"):                  # how much of the measured link the host path uses
            blk = line.get(key)
            if isinstance(blk, dict) and blk.get("ms_per_step") and host_link.get("h2d_gbs"):
                blk["h2d_gbs"] = blk["h2d_bytes_per_step"] * world / (blk["ms_per_step"] * 1e-3) / 1e9
                blk["link_frac"] = blk["h2d_gbs"] / host_link["h2d_gbs"]
    if long_signal is not None:
        line["long_signal"] = long_signal
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def probe_host_link(torch, dev, barrier, max_over_ranks, sum_over_ranks, mb=512, reps=4):
    """Aggregate pinned-copy bandwidth with every rank copying at once: H2D only, D2H only, both directions."""
    n = mb << 20
    h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    d_in = torch.empty(n, dtype=torch.uint8, device=dev)
    d_out = torch.empty(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    res = {}
    for mode in ("h2d", "d2h", "both"):
        for timed in (False, True):
            barrier()
            w0 = time.perf_counter()
            for _ in range(reps):
                if mode in ("h2d", "both"):
                    with torch.cuda.stream(s1):
                        d_in.copy_(h_in, non_blocking=True)
                if mode in ("d2h", "both"):
                    with torch.cuda.stream(s2):
                        h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()
            dt = max_over_ranks(time.perf_counter() - w0)
        per_dir = sum_over_ranks(float(n) * reps) / dt / 1e9
        res[mode + "_gbs"] = per_dir                         # per direction, summed over ranks
    res["what"] = "pinned cudaMemcpyAsync, %d MiB x %d per direction and rank, all ranks at once, GB/s summed " \
                  "over ranks (per direction)" % (mb, reps)
    return res


def run_long_signal(torch, dist, dev, local, world, rank, barrier, max_over_ranks, minutes=60, reps=5):
    """BASELINE config 4 on a shortened signal (SuperFlux, hop 441): contiguous frame ranges with a frame_size
    halo per rank, three ways of getting the full (T, 140) matrix onto every rank -- kernel + one NCCL all-gather
    (LongSignalFrontEnd.run), the kernel storing to all peers itself (run_fused, peer stores), the same through
    the NVLS multicast mapping -- compared bit for bit with each other and with the unsharded launch."""
    from madmom_b200 import parallel, recipe as R
    from madmom_b200.batch import BatchFrontEnd
    from madmom_b200.synth import make_batch_torch
    # rank 0 draws the signal and broadcasts it (the comparison below must not depend on the generator giving the
    # same samples on every device), then every rank keeps only its slice
    whole = make_batch_torch(minutes, 60.0, device=dev, seed=7).reshape(-1)
    if world > 1:
        dist.broadcast(whole.view(torch.uint8), src=0)        # NCCL has no int16: the bytes travel
    n = whole.numel()
    fe = parallel.LongSignalFrontEnd(R.SUPERFLUX, device=local)
    shard = fe.shard(n, world, rank)
    mine = whole[shard.sample_lo:shard.sample_hi].clone()
    ref = BatchFrontEnd(R.SUPERFLUX, np.int16, device=local).run_device(whole, [n], ("diff",))["diff"]
    torch.cuda.synchronize()
    del whole
    res = {"signal_minutes": minutes, "frames": int(ref.shape[0]), "ranks": world, "recipe": "SuperFlux (T, 140)",
           "samples_per_rank": int(shard.sample_hi - shard.sample_lo)}
    variants = {"nccl_allgather": lambda: fe.run(mine, n, ("diff",))["diff"],
                "fused_peer_stores": lambda: fe.run_fused(mine, n, ("diff",), multicast=False)["diff"]}
    results = {}
    for name, fn in list(variants.items()) + [("fused_multicast", None)]:
        try:
            if fn is None:
                fn = lambda: fe.run_fused(mine, n, ("diff",), multicast=True)["diff"]
            got = fn()
            torch.cuda.synchronize()
            for _ in range(2):
                fn()
            barrier()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for _ in range(reps):
                got = fn()
            t1.record()
            barrier()
            res[name + "_ms"] = max_over_ranks(t0.elapsed_time(t1)) / reps
            results[name] = got.clone()
        except Exception as exc:
            res[name + "_error"] = "%s: %s" % (type(exc).__name__, str(exc)[:200])
    names = ["fused_multicast", "fused_peer_stores", "nccl_allgather"]      # fixed shape: the same on every rank
    same = [-1.0 if k not in results else
            1.0 if results[k].shape == ref.shape and torch.equal(results[k], ref) else 0.0 for k in names]
    flag = torch.tensor(same, dtype=torch.float64, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)          # a variant counts only if it ran and matched on every rank
    flag = dict(zip(names, flag.tolist()))
    res["differs"] = [k for k in names if flag[k] == 0.0]
    names = [k for k in names if flag[k] >= 0.0]
    res["bit_identical"] = not res["differs"] and len(names) >= 2
    res["compared"] = names + ["unsharded launch"]
    return res


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
    ap.add_argument("--no-e2e", action="store_true", help="skip the host end-to-end legs (profiling runs)")
    ap.add_argument("--no-configs", action="store_true", help="skip the timings of the other configurations")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
