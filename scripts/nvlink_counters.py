#!/usr/bin/env python
"""NVLink data counters around the three ways of gathering the long-signal result (BASELINE config 4):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 \
        scripts/nvlink_counters.py [--hours 10] [--steps 20]

For each variant -- kernel + NCCL all-gather (LongSignalFrontEnd.run), the kernel storing to all peers
(mmb_run_gather), the kernel storing through the NVLS multicast mapping (mmb_run_multicast) -- rank 0 reads
`nvidia-smi nvlink -gt d` (data Tx / Rx KiB per link, all GPUs) before and after `steps` back-to-back steps between
barriers and prints one JSON line: ms per step (CUDA events, max over ranks), NVLink bytes per step and GPU in each
direction (sum over the GPU's links, mean / min / max over GPUs), the payload the variant has to move per rank
(rows of the other ranks: (R - 1) / R of the (T, 160-float padded) matrix in, the own slice out once per peer or once
in all), and the receive rate that follows from counter bytes and time.  The counters count payload bytes of the
link's data packets; whatever the driver itself moves in the meantime (barrier flags) is in there too."""
import argparse
import json
import os
import re
import subprocess
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import parallel, recipe as R           # noqa: E402
from madmom_b200.synth import make_batch_torch          # noqa: E402


def read_counters():
    """{gpu index: (tx_bytes, rx_bytes)} summed over links, and the raw text."""
    raw = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d"], capture_output=True, text=True).stdout
    out, gpu = {}, None
    for line in raw.splitlines():
        m = re.match(r"\s*GPU (\d+):", line)
        if m:
            gpu = int(m.group(1))
            out[gpu] = [0, 0]
            continue
        m = re.search(r"Data (Tx|Rx):\s*(\d+)\s*KiB", line)
        if m and gpu is not None:
            out[gpu][0 if m.group(1) == "Tx" else 1] += int(m.group(2)) * 1024
    return out, raw


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hours", type=float, default=10.0)
    ap.add_argument("--steps", type=int, default=20)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n = int(round(args.hours * 3600 * 44100))
    fe = parallel.LongSignalFrontEnd(R.SUPERFLUX, np.int16, device=local)
    sh = fe.shard(n, world, rank)
    m = sh.sample_hi - sh.sample_lo
    sl = make_batch_torch(1, m / 44100.0, device=dev, seed=7 + rank).reshape(-1)[:m].contiguous()
    T = R.num_frames(n, R.SUPERFLUX.hop_size)
    cols = fe.plan.width("diff")
    padded = -(-cols // 32) * 32
    variants = (("nccl_allgather", lambda: fe.run(sl, n, ("diff",))["diff"], cols),
                ("fused_peer_stores", lambda: fe.run_fused(sl, n, ("diff",), multicast=False)["diff"], cols),
                ("fused_multicast", lambda: fe.run_fused(sl, n, ("diff",), multicast=True)["diff"], cols))
    for name, fn, width in variants:
        try:
            for _ in range(3):
                fn()
            barrier()
            before = read_counters() if rank == 0 else None
            barrier()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for _ in range(args.steps):
                fn()
            t1.record()
            barrier()
            ms = reduce_max(t0.elapsed_time(t1)) / args.steps
            after = read_counters() if rank == 0 else None
            barrier()
        except Exception as exc:
            if rank == 0:
                print(json.dumps({"variant": name, "error": "%s: %s" % (type(exc).__name__, str(exc)[:300])}), flush=True)
            continue
        if rank != 0:
            continue
        (b, _), (a, raw) = before, after
        line = {"variant": name, "ranks": world, "hours": args.hours, "frames": T, "steps": args.steps,
                "ms_per_step": ms, "rows_per_rank": sh.rows_per_rank, "cols": cols}
        own = sh.rows_per_rank * width * 4
        line["payload_bytes_per_rank"] = {"rx": own * (world - 1),
                                          "tx": own * (1 if name == "fused_multicast" else world - 1)}
        if a and set(a) == set(b) and len(a) >= world:
            tx = np.array([(a[g][0] - b[g][0]) / args.steps for g in sorted(a)[:world]], dtype=np.float64)
            rx = np.array([(a[g][1] - b[g][1]) / args.steps for g in sorted(a)[:world]], dtype=np.float64)
            line["nvlink_bytes_per_step_per_gpu"] = {
                "tx_mean": tx.mean(), "tx_min": tx.min(), "tx_max": tx.max(),
                "rx_mean": rx.mean(), "rx_min": rx.min(), "rx_max": rx.max()}
            line["rx_gb_per_s_per_gpu"] = rx.mean() / (ms * 1e-3) / 1e9
            line["tx_gb_per_s_per_gpu"] = tx.mean() / (ms * 1e-3) / 1e9
        else:
            line["nvlink_counters"] = "not parsed"
            line["raw_head"] = raw[:1500]
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
