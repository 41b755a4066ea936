#!/usr/bin/env python
"""Device-resident time of each recipe of the multi-resolution stack on its own (1024 x 30 s tracks), with the plan's
warps / shared memory:   python scripts/per_n_probe.py [beat|onset]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import recipe as R
from madmom_b200.batch import BatchFrontEnd
from madmom_b200.synth import make_batch_torch

which = sys.argv[1] if len(sys.argv) > 1 else "beat"
recs = R.RNN_BEAT if which == "beat" else R.RNN_ONSET
pcm = make_batch_torch(1024, 30.0, device="cuda", seed=1000)
lengths = [pcm.shape[1]] * 1024
for rec in recs:
    fe = BatchFrontEnd([rec])
    units, T = fe.units(lengths)
    rows = int(sum(T))
    out = fe.run_device(pcm, lengths, ("log", "diff"), units=units, rows=rows)
    for _ in range(3):
        fe.run_device(pcm, lengths, ("log", "diff"), out=out, units=units, rows=rows)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        fe.run_device(pcm, lengths, ("log", "diff"), out=out, units=units, rows=rows)
    b.record()
    torch.cuda.synchronize()
    info = fe.plans[0].info
    print("N=%d MMB_WARPS=%s ms %.3f warps %d fpr %d smem %d bands %d" % (
        rec.frame_size, os.environ.get("MMB_WARPS", "-"), a.elapsed_time(b) / 5, info["warps"],
        info["frames_per_round"], info["smem_bytes"], fe.plans[0].num_bands))
