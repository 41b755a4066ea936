#!/usr/bin/env python
"""Time mmb_complex_flux alone on a matrix of config-5 size (rows x 1024 lgd values, 24-bands-per-octave bank of the
ComplexFlux recipe, 3000-row signals), once per kernel behind the entry point:

    python scripts/complex_flux_probe.py [rows]

cover   = complex_flux_cover_kernel (default), stream = complex_flux_stream_kernel (MMB_CF_STREAM=1),
generic = complex_flux_kernel (MMB_CF_GENERIC=1).  Prints ms per call, rows/s, GB/s of lgd read and whether the
results are bit-identical."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import _lib, recipe as R           # noqa: E402
from madmom_b200.batch import BatchFrontEnd         # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1536000
dev = torch.device("cuda:0")
fe = BatchFrontEnd(R.COMPLEXFLUX, np.int16, device=0)
plan = fe.plans[0]
K, B = plan.num_bins, plan.num_bands
fb = fe.recipes[0].build_filterbank()                     # the ranges complex_flux_batch passes: band +- one bin
starts, lens, _ = fb.band_table()
lo = torch.from_numpy(np.maximum(starts.astype(np.int64) - 1, 0).astype(np.int32)).to(dev)
hi = torch.from_numpy(np.minimum(starts.astype(np.int64) + lens + 1, fb.shape[0]).astype(np.int32)).to(dev)
g = torch.Generator(device=dev)
g.manual_seed(5)
lgd = torch.rand((rows, K), device=dev, generator=g)
diff = torch.rand((rows, B), device=dev, generator=g)
print("rows %d  bins %d  bands %d  widest band %d bins" % (rows, K, B, int((hi - lo).max())))
results = {}
for name, env in (("cover", None), ("stream", "MMB_CF_STREAM"), ("generic", "MMB_CF_GENERIC")):
    if env:
        os.environ[env] = "1"
    out = _lib.complex_flux(lgd, diff, lo, hi, 3, rows_per_signal=3000)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    t0.record()
    for _ in range(reps):
        out = _lib.complex_flux(lgd, diff, lo, hi, 3, rows_per_signal=3000)
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / reps
    results[name] = out.clone()
    print("%-8s %8.3f ms   %7.1f M rows/s   %6.0f GB/s of lgd" % (name, ms, rows / ms / 1e3, rows * K * 4 / ms / 1e6))
    if env:
        del os.environ[env]
print("bit-identical: cover == stream %s, cover == generic %s" % (
    torch.equal(results["cover"], results["stream"]), torch.equal(results["cover"], results["generic"])))
