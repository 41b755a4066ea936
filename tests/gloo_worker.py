"""Worker of tests/test_parallel_gloo.py: one rank of a world_size-N gloo group on CPU.  The per-rank compute is
the CPU oracle applied to exactly the PCM slice `madmom_b200.parallel.frame_shards` hands that rank (zeros
everywhere else), so a halo that is too small, a wrong global frame index or a wrong gather order all show up as a
mismatch with the unsharded oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world, port, hop = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], float(sys.argv[4])
    import torch
    import torch.distributed as dist
    from madmom_b200 import parallel, recipe as R
    from madmom_b200.synth import make_track
    from oracle import frontend as O
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % port, rank=rank, world_size=world)
    x = make_track(7, num_samples=150001)
    rec = dict(O.SUPERFLUX, hop_size=hop)
    d = O.diff_frames(0.5, hop, 2048)
    T = O.num_frames(len(x), hop)
    sh = parallel.frame_shards(len(x), hop, 2048, world, 0, d)[rank]
    local = np.zeros((sh.rows_per_rank, 140), np.float32)
    if sh.frame_count:
        z = np.zeros_like(x)
        z[sh.sample_lo:sh.sample_hi] = x[sh.sample_lo:sh.sample_hi]       # the rank sees only its slice
        lead = min(d, sh.frame_first)
        rows = O.run_recipe(z, rec, first_frame=sh.frame_first - lead, n_frames=sh.frame_count + lead)
        if lead == 0:
            local[:sh.frame_count] = rows
        else:
            local[:sh.frame_count] = rows[lead:]
    full = parallel.gather_rows(torch.from_numpy(local), T).numpy()
    lo, hi = parallel.shard_tracks(10, world, rank)
    counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([hi - lo]))
    if rank == 0:
        want = O.run_recipe(x, rec)
        ok = full.shape == want.shape and np.array_equal(full, want) and int(sum(c.item() for c in counts)) == 10
        print("OK" if ok else "MISMATCH max %g" % np.abs(full - want).max())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
