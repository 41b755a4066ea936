"""Multi-rank host logic on CPU (gloo, world_size 2 and 3): frame-range sharding with halos + one all-gather
reproduces the unsharded result bit for bit; track sharding covers every track exactly once."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,hop", [(2, 441.), (3, 220.5)])
def test_frame_range_sharding_gloo(world, hop):
    port = str(_free_port())
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "gloo_worker.py"), str(r), str(world),
                               port, str(hop)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
             for r in range(world)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "OK" in outs[0], outs[0]


def test_shard_arithmetic():
    from madmom_b200 import parallel, recipe as R
    for n, w in ((10, 3), (8192, 8), (5, 8), (0, 2)):
        blocks = [parallel.shard_tracks(n, w, r) for r in range(w)]
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
        assert max(b[1] - b[0] for b in blocks) - min(b[1] - b[0] for b in blocks) <= 1
    # the 10-hour signal of BASELINE config 4 on 8 ranks: equal row counts, halos of N/2 + d*hop / N/2 samples
    n = 1587600000
    sh = parallel.frame_shards(n, 441., 2048, 8, 0, 1)
    assert sum(s.frame_count for s in sh) == R.num_frames(n, 441.) == 3600000
    assert len({s.rows_per_rank for s in sh}) == 1 and sh[0].rows_per_rank == 450000
    assert sh[0].sample_lo == 0 and sh[-1].sample_hi == n
    for s in sh[1:]:
        assert s.sample_lo == (s.frame_first - 1) * 441 - 1024
    for s in sh[:-1]:
        assert s.sample_hi == (s.frame_first + s.frame_count - 1) * 441 + 1024
    assert s.unit() == (0, s.sample_hi - s.sample_lo, s.sample_lo, s.frame_first, s.frame_count, 0)
    # more ranks than frames: trailing ranks are empty but still contribute padded rows
    sh = parallel.frame_shards(1000, 441., 2048, 4, 0, 1)
    assert [s.frame_count for s in sh] == [1, 1, 1, 0]


def test_frame_shards_for_paired_frames():
    """frame_size 1024: the kernel packs frames (2m, 2m + 1) into one FFT, so every range starts at an even frame
    (range size rounded up to even) and the left halo reaches one frame further back than diff_frames alone needs;
    the ranges still cover every frame exactly once."""
    from madmom_b200 import parallel, recipe as R
    n, hop, d = 44100 * 61 + 123, 441., 1
    T = R.num_frames(n, hop)
    for world in (2, 3, 5, 8):
        sh = parallel.frame_shards(n, hop, 1024, world, 0, d)
        per = sh[0].rows_per_rank
        assert per % 2 == 0 and per * world >= T and (per - 2) * world < T + 2 * world
        assert sum(s.frame_count for s in sh) == T
        at = 0
        for s in sh:
            if s.frame_count == 0:
                continue
            assert s.frame_first == at and s.frame_first % 2 == 0
            at += s.frame_count
            lead = max(s.frame_first - d - 1, 0)
            assert s.sample_lo == max(R.frame_start(lead, 1024, hop, 0), 0)
            assert s.sample_hi == min(R.frame_start(s.frame_first + s.frame_count - 1, 1024, hop, 0) + 1024, n)
        # the other frame sizes keep the plain ceil(T / world) split
        assert parallel.frame_shards(n, hop, 2048, world, 0, d)[0].rows_per_rank == -(-T // world)
