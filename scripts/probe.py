"""Quick device-side throughput probe (not the bench): frames/s of the fused kernel for a few recipes, int16 and
float32 PCM, through the package's own recipe layer (nothing from oracle/).

    python scripts/probe.py [tracks]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import _lib, recipe as R               # noqa: E402
from madmom_b200.batch import BatchFrontEnd             # noqa: E402
from madmom_b200.synth import make_batch_torch          # noqa: E402


def probe(name, rec, n_tracks, dtype, outs=("diff", "odf"), seconds=30.0, reps=5):
    pcm = make_batch_torch(n_tracks, seconds, dtype=None if dtype == np.int16 else torch.float32)
    lengths = [pcm.shape[1]] * n_tracks
    fe = BatchFrontEnd(rec, dtype, device=0)
    out = fe.run_device(pcm, lengths, outs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fe.run_device(pcm, lengths, outs, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    rows = out[outs[0]].shape[0]
    print("%-16s %-7s outs=%-9s %8.3f ms  %7.1f M frames/s  %6.2f M x realtime" % (
        name, np.dtype(dtype).name, ",".join(outs), ms, rows / ms / 1e3, n_tracks * seconds / ms / 1e3), flush=True)


if __name__ == "__main__":
    tracks = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    print("fp32 peak TFLOP/s:", _lib.measure_fp32_peak(0), flush=True)
    for name, rec in (("LogFiltSpecFlux", R.LOGFILTSPECFLUX), ("SuperFlux", R.SUPERFLUX)):
        for dtype in (np.int16, np.float32):
            probe(name, rec, tracks, dtype)
    for N in (1024, 4096):
        rec = R.LOGFILTSPECFLUX.replace(frame_size=N)
        for dtype in (np.int16, np.float32):
            probe("LogFilt N=%d" % N, rec, tracks, dtype)
