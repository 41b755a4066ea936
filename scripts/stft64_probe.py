#!/usr/bin/env python
"""Time the float64 STFT kernel (csrc/stft64.cu) that serves the raw `stft` / `spec` outputs, beside the fused FP32
kernel on the same tracks:   python scripts/stft64_probe.py [tracks]   (30 s tracks, frame size 2048, hop 441)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import recipe as R                    # noqa: E402
from madmom_b200.batch import BatchFrontEnd            # noqa: E402
from madmom_b200.synth import make_batch_torch         # noqa: E402

tracks = int(sys.argv[1]) if len(sys.argv) > 1 else 128
pcm = make_batch_torch(tracks, 30.0, device="cuda", seed=1000)
lengths = [pcm.shape[1]] * tracks
fe = BatchFrontEnd(R.LOGFILTSPECFLUX, np.int16, device=0)
for outputs in (("diff", "odf"), ("spec",), ("stft",), ("stft", "spec")):
    out = fe.run_device(pcm, lengths, outputs)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(5):
        fe.run_device(pcm, lengths, outputs, out=out)
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / 5
    frames = out[outputs[0]].shape[0]
    nbytes = sum(t.numel() * t.element_size() for t in out.values())
    print("%-18s %8.3f ms  %7.1f M frames/s  %8.0f x realtime  outputs %6.2f GB -> %5.0f GB/s written" % (
        "+".join(outputs), ms, frames / ms / 1e3, tracks * 30.0 / (ms * 1e-3), nbytes / 1e9, nbytes / ms / 1e6))
    del out
