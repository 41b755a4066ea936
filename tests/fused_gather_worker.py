"""Worker of tests/test_gpu_fused_gather.py (one process per GPU under torchrun): the fused front end + all-gather
(LongSignalFrontEnd.run_fused, mmb_run_gather over torch symmetric memory) must equal the NCCL path
(LongSignalFrontEnd.run: kernel, then one all_gather_into_tensor) bit for bit on every rank."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from madmom_b200 import parallel, recipe as R
    from madmom_b200.synth import make_track
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    x = make_track(21, num_samples=44100 * 61 + 123)          # every rank builds the same signal
    for rec in (R.SUPERFLUX, R.SUPERFLUX_200, R.RNN_ONSET[0]):
        fe = parallel.LongSignalFrontEnd(rec)
        sh = fe.shard(len(x), world, rank)
        pcm = torch.from_numpy(x[sh.sample_lo:sh.sample_hi].copy()).cuda()
        want = fe.run(pcm, len(x), ("log", "diff", "odf"))
        for mc in (False, None, False, None):                  # peer stores / NVLS multicast when the box has it;
            got = fe.run_fused(pcm, len(x), ("log", "diff", "odf"), multicast=mc)   # the buffers are reused
            torch.cuda.synchronize()
            for k in ("log", "diff", "odf"):
                assert got[k].shape == want[k].shape, (k, got[k].shape, want[k].shape)
                assert torch.equal(got[k], want[k]), (rec.frame_size, rec.hop_size, k, rank)
    dist.barrier()
    if rank == 0:
        print("fused gather ok on %d ranks" % world, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
