#!/usr/bin/env python
"""Launches each kernel variant the profiles/ summaries cover twice (256 tracks x 30 s): the three frame sizes of the
multi-resolution stack, the SuperFlux recipe, and the ComplexFlux pair (front end with lgd output + streaming
mmb_complex_flux).  Run under ncu with -k regex:'frontend_kernel|complex_flux'."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import recipe as R
from madmom_b200.batch import BatchFrontEnd, complex_flux_batch
from madmom_b200.synth import make_batch_torch

tracks = int(sys.argv[1]) if len(sys.argv) > 1 else 256
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
pcm = make_batch_torch(tracks, 30.0, device="cuda", seed=1000)
lengths = [pcm.shape[1]] * tracks
for rec in list(R.RNN_BEAT) + [R.SUPERFLUX]:
    fe = BatchFrontEnd([rec])
    for _ in range(reps):
        out = fe.run_device(pcm, lengths, ("log", "diff"))
    torch.cuda.synchronize()
    print("N=%d bands %d frames %d" % (rec.frame_size, fe.plans[0].num_bands, out["diff"].shape[0]))
fe = BatchFrontEnd([R.COMPLEXFLUX])
for _ in range(reps):
    odf = complex_flux_batch(fe, pcm, lengths)
torch.cuda.synchronize()
print("complexflux frames %d" % odf.shape[0])
