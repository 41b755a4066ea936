#!/usr/bin/env python
"""The host's cut of a launch into the fused kernel's work list (csrc/work_list.h), measured on batches of
different shapes: the balanced list (default) against equal chunks of 32 and 64 rounds dealt in turn (a plan with
`chunk_frames` set: the design before it).

    python scripts/work_list_probe.py [--reps 10]

One JSON line per batch: tracks, audio seconds, ms per launch (CUDA events, L2 flushed by the batch itself where it
is larger than L2) for the three cuts, and that the three give bit-identical rows."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import recipe as R                      # noqa: E402
from madmom_b200.batch import BatchFrontEnd              # noqa: E402
from madmom_b200.synth import make_batch_torch           # noqa: E402

SR = 44100


def batches():
    rng = np.random.default_rng(11)
    yield "1024 x 30 s (config 2)", [30.0] * 1024
    yield "200 x 30 s", [30.0] * 200
    yield "64 x 30 s", [30.0] * 64
    yield "1024 x uniform 5..60 s", rng.uniform(5., 60., 1024).tolist()
    yield "256 x uniform 5..60 s", rng.uniform(5., 60., 256).tolist()
    yield "512 x log-normal, median 30 s", np.minimum(30. * rng.lognormal(0., 0.7, 512), 240.).tolist()
    yield "16 x 10 min", [600.0] * 16


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    rec = R.LOGFILTSPECFLUX
    fpr = 16                                                # frames per round, N = 2048
    cuts = (("balanced", 0), ("chunks32", 32 * fpr - 1), ("chunks64", 64 * fpr - 1))
    fes = {name: BatchFrontEnd(rec, np.int16, device=0, chunk_frames=c) for name, c in cuts}
    base = make_batch_torch(64, 30.0, device="cuda:0", seed=3).reshape(-1)
    for label, seconds in batches():
        lengths = [int(round(s * SR)) for s in seconds]
        total = int(sum(lengths))
        # 64 synthetic tracks repeated and cut into the batch's tracks (the content does not matter for the timing)
        pcm = base.repeat(-(-total // base.numel()))[:total].contiguous()
        line = {"batch": label, "tracks": len(lengths), "audio_s": total / float(SR)}
        ref = None
        for name, _ in cuts:
            fe = fes[name]
            out = fe.run_device(pcm, lengths, ("diff", "odf"))
            for _ in range(2):
                fe.run_device(pcm, lengths, ("diff", "odf"), out=out)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.reps):
                fe.run_device(pcm, lengths, ("diff", "odf"), out=out)
            e1.record()
            torch.cuda.synchronize()
            line["ms_" + name] = e0.elapsed_time(e1) / args.reps
            if ref is None:
                ref = {k: v.clone() for k, v in out.items()}
            else:
                line["same_rows_" + name] = all(torch.equal(ref[k].view(torch.int32), out[k].view(torch.int32))
                                                for k in ref)
            del out
        line["balanced_vs_chunks32"] = line["ms_chunks32"] / line["ms_balanced"]
        line["balanced_vs_chunks64"] = line["ms_chunks64"] / line["ms_balanced"]
        print(json.dumps(line), flush=True)
        del pcm, ref
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
