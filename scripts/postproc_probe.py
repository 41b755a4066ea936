#!/usr/bin/env python
"""Achieved bandwidth of the post-processing kernels (csrc/postproc.cu) against the bytes each one must move:

    python scripts/postproc_probe.py [rows]          (default 384 000 rows = 128 tracks x 30 s at 100 fps)

One line per kernel: ms per call, the algorithmic bytes (inputs read once + outputs written once), GB/s, and the
fraction of the measured HBM copy bandwidth (MEASURED_PEAKS.json, else 6500 GB/s)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from madmom_b200 import _lib        # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 384000
dev = torch.device("cuda:0")
peak = 6500.0
try:
    mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    peak = float(mp.get("hbm_gbs", mp.get("hbm_copy_gbs", peak)))
except Exception:
    pass
g = torch.Generator(device=dev)
g.manual_seed(1)
K, B = 1024, 140
stft = torch.view_as_complex(torch.randn((rows, K, 2), device=dev, generator=g))
spec = stft.abs().contiguous()
logf = torch.rand((rows, B), device=dev, generator=g)
act = (torch.rand((rows,), device=dev, generator=g) ** 4) * 3
seg = torch.arange(0, rows + 1, 3000, device=dev, dtype=torch.int64)
if int(seg[-1]) != rows:
    seg = torch.cat((seg, torch.tensor([rows], device=dev, dtype=torch.int64)))
pcm2 = (torch.randn((rows * 441, 2), device=dev, generator=g) * 3000).to(torch.int16)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(reps):
        fn()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / reps


cases = [
    ("matrix_diff (rows x 1024, d=1, positive)", lambda: _lib.matrix_diff(spec, 1, 0, True, rows_per_signal=3000),
     2 * rows * K * 4),
    ("matrix_diff (rows x 1024, max_bins=3)", lambda: _lib.matrix_diff(spec, 1, 3, True, rows_per_signal=3000),
     2 * rows * K * 4),
    ("matrix_diff flux only (rows x 140)", lambda: _lib.matrix_diff(logf, 1, 3, True, rows_per_signal=3000,
                                                                    want_diff=False, want_odf=True), rows * B * 4),
    ("phase_lgd: phase", lambda: _lib.phase_lgd(stft, True, False), rows * K * (8 + 4)),
    ("phase_lgd: lgd", lambda: _lib.phase_lgd(stft, False, True), rows * K * (8 + 4)),
    ("phase_lgd: phase + lgd", lambda: _lib.phase_lgd(stft, True, True), rows * K * (8 + 8)),
    ("peak_picking (SuperFlux parameters)", lambda: _lib.peak_picking(act, 1.1, 0, 30, 0, 2, 10, seg), rows * 8),
    ("remix_mono (stereo int16)", lambda: _lib.remix_mono(pcm2), pcm2.numel() * 2 + pcm2.shape[0] * 2),
]
for mode in _lib.ODF_MODES:
    m = _lib.ODF_MODES[mode]
    if m <= 3:
        cases.append(("spectral_odf %s" % mode, (lambda mode=mode: _lib.spectral_odf(mode, spec=spec, rows_per_signal=3000)),
                      rows * K * 4))
    else:
        cases.append(("spectral_odf %s" % mode, (lambda mode=mode: _lib.spectral_odf(mode, stft=stft, rows_per_signal=3000)),
                      rows * K * 8))
print("rows %d, HBM peak %.0f GB/s" % (rows, peak))
for name, fn, nbytes in cases:
    try:
        ms = timed(fn)
        gbs = nbytes / ms / 1e6
        print("%-52s %8.3f ms  %6.2f GB  %7.0f GB/s  %5.2f of HBM" % (name, ms, nbytes / 1e9, gbs, gbs / peak))
    except Exception as exc:
        print("%-52s failed: %s: %s" % (name, type(exc).__name__, str(exc)[:120]))
