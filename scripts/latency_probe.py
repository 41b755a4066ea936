#!/usr/bin/env python
"""Latency of the drop-in path for ONE file (BASELINE config 1: SuperFlux on sample.wav, 2.8 s, 281 frames):
what a madmom user sees who swaps the import and keeps processing files one at a time.

    python scripts/latency_probe.py [--iters 300] [--profile]

Prints one JSON line per case: median / p10 / p90 wall time per call in microseconds, host PCM in, host result out
(the call ends with np.asarray(result), i.e. includes H2D, the launch and D2H).  `--profile` adds a cProfile
listing of the first case on stderr.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def measure(fn, iters, warmup=20):
    import torch
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    ts = np.sort(np.asarray(ts)) * 1e6
    return {"median_us": float(np.median(ts)), "p10_us": float(ts[len(ts) // 10]),
            "p90_us": float(ts[(9 * len(ts)) // 10])}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=300)
    ap.add_argument("--profile", action="store_true")
    args = ap.parse_args()
    import torch
    from madmom_b200.audio.signal import Signal, FramedSignalProcessor
    from madmom_b200.audio.spectrogram import SuperFluxProcessor
    from madmom_b200.audio.filters import LogarithmicFilterbank
    from madmom_b200.features.onsets import SpectralOnsetProcessor, OnsetPeakPickingProcessor
    from madmom_b200.processors import SequentialProcessor

    z = np.load(os.path.join(ROOT, "tests", "golden", "sample_wav.npz"))
    pcm = np.ascontiguousarray(z["pcm"])
    sr = int(z["sample_rate"])
    long_pcm = np.tile(pcm, 64)[: 30 * sr * 6]                       # a 3-minute file
    torch.cuda.set_device(0)

    frames = FramedSignalProcessor(frame_size=2048, fps=200)
    superflux = SequentialProcessor([frames, SuperFluxProcessor()])
    # the parameter set of the reference's bin/SuperFlux program (bin/SuperFlux:59-69)
    odf = SpectralOnsetProcessor(onset_method="superflux", fps=200, frame_size=2048,
                                 filterbank=LogarithmicFilterbank, num_bands=24,
                                 norm_filters=False, log=np.log10, mul=1, add=1, diff_ratio=0.5, diff_max_bins=3,
                                 positive_diffs=True)
    pick = OnsetPeakPickingProcessor(threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                                     combine=0.03, delay=0, fps=200)
    program = SequentialProcessor([odf, pick])

    cases = [
        ("superflux_features_sample_wav", lambda: np.asarray(superflux(Signal(pcm, sample_rate=sr))), len(pcm)),
        ("superflux_program_onsets_sample_wav", lambda: np.asarray(program(Signal(pcm, sample_rate=sr))), len(pcm)),
        ("superflux_program_onsets_3min", lambda: np.asarray(program(Signal(long_pcm, sample_rate=sr))),
         len(long_pcm)),
    ]
    for name, fn, n in cases:
        r = measure(fn, args.iters)
        r.update({"case": name, "audio_s": n / sr, "x_realtime": n / sr / (r["median_us"] * 1e-6)})
        print(json.dumps(r), flush=True)
    if args.profile:
        import cProfile
        import pstats
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(200):
            cases[1][1]()
        pr.disable()
        pstats.Stats(pr, stream=sys.stderr).sort_stats("cumulative").print_stats(45)


if __name__ == "__main__":
    main()
