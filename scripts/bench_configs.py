#!/usr/bin/env python
"""Throughput of the other BASELINE.json configurations (bench.py measures configs[1]); one JSON line each.

    python scripts/bench_configs.py [--tracks 1024] [--steps 5] [--configs superflux,multires_beat,complexflux,long]
    torchrun --nproc-per-node N scripts/bench_configs.py --configs long      # config 4: frame ranges + one all-gather

Inputs are synthetic (madmom_b200.synth) and resident in HBM before the timed region; timing with CUDA events,
max over ranks.  audio-seconds/second = frames / fps / time.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(step, steps, warmup, torch, barrier):
    for _ in range(warmup):
        step()
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step()
    b.record()
    barrier()
    return a.elapsed_time(b) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tracks", type=int, default=1024)
    ap.add_argument("--seconds", type=float, default=30.0)
    ap.add_argument("--hours", type=float, default=10.0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--configs", default="superflux,multires_beat,multires_onset,complexflux,long")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from madmom_b200 import _lib, parallel, recipe as R
    from madmom_b200.batch import BatchFrontEnd, complex_flux_batch
    from madmom_b200.synth import make_batch_torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    wanted = args.configs.split(",")
    pcm = None
    if any(c != "long" for c in wanted):
        pcm = make_batch_torch(args.tracks, args.seconds, device=dev, seed=1000 + rank)
        lengths = [pcm.shape[1]] * args.tracks
    for name in wanted:
        l0 = _lib.launch_count
        if name in ("superflux", "multires_beat", "multires_onset"):
            recs, outs = {"superflux": ([R.SUPERFLUX], ("diff",)), "multires_beat": (list(R.RNN_BEAT), ("stack",)),
                          "multires_onset": (list(R.RNN_ONSET), ("stack",))}[name]
            fe = BatchFrontEnd(recs, np.int16, device=local)
            units, T = fe.units(lengths)
            rows = int(sum(T))
            out = fe.run_device(pcm, lengths, outs, units=units, rows=rows)
            ms = reduce_max(timed(lambda: fe.run_device(pcm, lengths, outs, out=out, units=units, rows=rows),
                                  args.steps, args.warmup, torch, barrier))
            fps, frames = recs[0].fps, rows * world
            cfg = {"workload": name, "tracks_per_gpu": args.tracks, "seconds": args.seconds,
                   "outputs": {k: list(v.shape) for k, v in out.items()}}
            del out
        elif name == "complexflux":
            fe = BatchFrontEnd([R.COMPLEXFLUX], np.int16, device=local)
            rows = int(sum(fe.frame_counts(lengths)))
            ms = reduce_max(timed(lambda: complex_flux_batch(fe, pcm, lengths), args.steps, args.warmup, torch, barrier))
            fps, frames = R.COMPLEXFLUX.fps, rows * world
            cfg = {"workload": name, "tracks_per_gpu": args.tracks, "seconds": args.seconds, "outputs": {"odf": [rows]},
                   "note": "lgd intermediate (rows x 1024 f32) goes through HBM in groups of <= 2 GiB"}
        elif name == "long":
            # config 4: one signal of `hours`, split into frame ranges with halos, one all-gather of the slices
            n = int(round(args.hours * 3600 * 44100))
            fe = parallel.LongSignalFrontEnd(R.SUPERFLUX, np.int16, device=local)
            sh = fe.shard(n, world, rank)
            m = sh.sample_hi - sh.sample_lo
            sl = make_batch_torch(1, m / 44100.0, device=dev, seed=7 + rank).reshape(-1)[:m].contiguous()
            T = R.num_frames(n, R.SUPERFLUX.hop_size)

            def step():
                loc = fe.run_local(sl, sh, ("diff",))
                return parallel.gather_rows(loc["diff"], T) if world > 1 else loc["diff"]
            ms = reduce_max(timed(step, args.steps, args.warmup, torch, barrier))
            fps, frames = R.SUPERFLUX.fps, T
            cfg = {"workload": "long_signal_superflux", "hours": args.hours, "ranks": world,
                   "rows_per_rank": sh.rows_per_rank, "halo_samples": [1024 + fe.plan.diff_frames * 441, 1024],
                   "collective": "one all_gather_into_tensor of (rows_per_rank, 140) f32 per rank" if world > 1
                   else "none (single rank)"}
            if world > 1:
                # the same step with the all-gather fused into the kernel (rows stored straight into every rank's
                # result over NVLink, mmb_run_gather), checked against the NCCL result, then timed
                for mc, how in ((False, "none: kernel stores each row to all ranks' symmetric-memory buffers "
                                 "(mmb_run_gather) + one barrier"),
                                (True, "none: kernel stores each row once through the NVLS multicast mapping of the "
                                 "ranks' symmetric-memory buffers (mmb_run_multicast) + one barrier")):
                    try:
                        got = fe.run_fused(sl, n, ("diff",), multicast=mc)["diff"]
                    except RuntimeError as e:                     # no multicast mapping on this box
                        if rank == 0:
                            print(json.dumps({"config": dict(cfg, collective=how), "unavailable": str(e)}), flush=True)
                        continue
                    torch.cuda.synchronize()
                    same = bool(torch.equal(got, step()))
                    del got
                    ms_f = reduce_max(timed(lambda: fe.run_fused(sl, n, ("diff",), multicast=mc), args.steps,
                                            args.warmup, torch, barrier))
                    if rank == 0:
                        c2 = dict(cfg, collective=how, bit_identical_to_nccl_path=same)
                        print(json.dumps({"metric": "audio-seconds per second (x realtime)",
                                          "value": frames / fps / (ms_f * 1e-3), "unit": "audio-s/s", "n_gpus": world,
                                          "ms_per_step": ms_f, "steps": args.steps, "frames_per_step": frames,
                                          "config": c2}), flush=True)
            del sl
        else:
            raise SystemExit("unknown config %r" % name)
        if rank == 0:
            print(json.dumps({"metric": "audio-seconds per second (x realtime)", "value": frames / fps / (ms * 1e-3),
                              "unit": "audio-s/s", "n_gpus": world, "ms_per_step": ms, "steps": args.steps,
                              "frames_per_step": frames, "gpu_launches_per_step":
                              (_lib.launch_count - l0) / float(args.steps + args.warmup + 1), "config": cfg}))
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
