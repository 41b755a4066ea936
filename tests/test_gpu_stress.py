"""Randomised small cases against the oracle: many short and empty tracks (chunk changes every round, partial rounds,
rounds that are all zero padding), odd offsets, fractional hops, every frame size, int16 and float32, lags up to the
ring depth.  The kernel has no CTA barrier in its steady state; a synchronisation mistake shows up here as a wrong
row."""
import numpy as np
import pytest

from conftest import assert_close
from test_gpu_plan import _plan_for, _run_all

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _tracks(rng, n_tracks, max_len, dtype):
    out = []
    for _ in range(n_tracks):
        n = int(rng.choice([1, 2, 7, 440, 441, 442, 1023, 2049, rng.integers(1, max_len)]))
        x = rng.integers(-20000, 20000, size=n).astype(np.int16)
        if n > 600:
            t = np.arange(n)
            x = (x // 8 + (6000 * np.sin(2 * np.pi * rng.uniform(100, 5000) * t / 44100.0)).astype(np.int16))
        out.append(x if dtype == np.int16 else (x / 32767.0).astype(np.float32))
    return out


@pytest.mark.parametrize("seed", range(12))
def test_random_batches_vs_oracle(seed):
    from oracle import frontend as O
    rng = np.random.default_rng(100 + seed)
    N = int(rng.choice([1024, 2048, 4096]))
    hop = float(rng.choice([441., 220.5, 512., 147.3]))
    dtype = np.int16 if rng.random() < 0.7 else np.float32
    bpo, norm = (6, True) if rng.random() < 0.5 else (24, False)
    ratio = float(rng.choice([0.5, 0.25]))
    max_bins = 3 if rng.random() < 0.5 else None
    chunk = int(rng.choice([0, 5, 33]))
    plan, recipe = _plan_for(dtype, N, hop, bpo, norm, ratio=ratio, max_bins=max_bins, chunk_frames=chunk)
    tracks = _tracks(rng, int(rng.integers(20, 60)), 9000, dtype)
    got, T = _run_all(plan, tracks, hop)
    r0 = 0
    for i, x in enumerate(tracks):
        ref = O.run_recipe(x, recipe, stages=True)
        rows = slice(r0, r0 + T[i])
        assert ref["diff"].shape[0] == T[i]
        what = "seed %d track %d (N %d hop %s len %d)" % (seed, i, N, hop, len(x))
        assert_close(got["log"][rows], ref["log"], what=what + " log")
        assert_close(got["diff"][rows], ref["diff"], what=what + " diff")
        assert_close(got["odf"][rows], ref["odf"], what=what + " odf")
        r0 += T[i]
    assert r0 == got["diff"].shape[0]
    # and again in one go with the outputs pre-filled with NaN: every row must have been written exactly as before
    again, _ = _run_all(plan, tracks, hop)
    for k in ("log", "diff", "odf"):
        assert np.array_equal(again[k], got[k]), k


def test_many_tiny_tracks():
    """20 000 tracks of 1 .. 3000 samples (1 .. 7 frames each): more chunks than the grid has CTAs many times over,
    every round is a chunk change.  Sampled tracks against the oracle, everything twice for reproducibility."""
    from oracle import frontend as O
    rng = np.random.default_rng(7)
    plan, recipe = _plan_for(np.int16, 2048, 441., 12, False)
    tracks = [rng.integers(-8000, 8000, size=int(n)).astype(np.int16) for n in rng.integers(1, 3000, size=20000)]
    got, T = _run_all(plan, tracks, 441.)
    again, _ = _run_all(plan, tracks, 441.)
    assert np.array_equal(got["diff"], again["diff"]) and np.array_equal(got["odf"], again["odf"])
    assert not np.isnan(got["diff"]).any() and not np.isnan(got["odf"]).any()
    starts = np.concatenate(([0], np.cumsum(T)))
    for i in rng.integers(0, len(tracks), size=60):
        ref = O.run_recipe(tracks[i], recipe, stages=True)
        rows = slice(starts[i], starts[i + 1])
        assert_close(got["diff"][rows], ref["diff"], what="tiny track %d" % i)
        assert_close(got["odf"][rows], ref["odf"], what="tiny track %d odf" % i)
