"""The host-side cut of a launch into the fused kernel's work list (madmom_b200/csrc/work_list.h), run on the CPU
through tests/host_emul.cpp: every frame of every unit exactly once, equal rounds for the warp groups, cuts of
1024-point units in front of even frames, the strided lay-out with its holes, fewer CTAs for a small launch."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "host_emul.cpp")
HDR = os.path.join(ROOT, "madmom_b200", "csrc", "work_list.h")
LIB = os.path.join(ROOT, "build", "libhost_emul.so")


@pytest.fixture(scope="module")
def emul():
    stale = not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(SRC), os.path.getmtime(HDR))
    if stale:
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-o", LIB, SRC])
    lib = ctypes.CDLL(LIB)
    i64p = ctypes.POINTER(ctypes.c_longlong)
    lib.emul_work_list.restype = ctypes.c_longlong
    lib.emul_work_list.argtypes = [ctypes.c_int] * 4 + [ctypes.c_longlong] * 3 + [i64p, ctypes.c_int, i64p,
                                                                                 ctypes.c_longlong,
                                                                                 ctypes.POINTER(ctypes.c_int)]
    return lib


def work_list(lib, N, fpr, groups, sms, lead, counts, firsts=None, fixed=0, wave=64):
    n = len(counts)
    units = np.zeros((n, 6), dtype=np.int64)
    units[:, 4] = counts
    if firsts is not None:
        units[:, 3] = firsts
    cap = 4 * (int(np.sum(counts)) // max(fpr, 1) + n + sms * groups) + 64
    out = np.zeros((cap, 3), dtype=np.int64)
    grid = ctypes.c_int(0)
    i64p = ctypes.POINTER(ctypes.c_longlong)
    m = lib.emul_work_list(N, fpr, groups, sms, lead, fixed, wave, units.ctypes.data_as(i64p), n,
                           out.ctypes.data_as(i64p), cap, ctypes.byref(grid))
    assert m >= 0
    return out[:m], grid.value


def check(work, grid, N, fpr, groups, sms, lead, counts, firsts=None, wave=64):
    counts = np.asarray(counts, dtype=np.int64)
    firsts = np.zeros_like(counts) if firsts is None else np.asarray(firsts, dtype=np.int64)
    total_frames = int(counts.sum())
    if total_frames == 0:
        assert len(work) == 0 or not work[:, 1].any()
        return None
    assert 1 <= grid <= sms
    stride = grid * groups
    assert len(work) % stride == 0
    # every frame of every unit exactly once
    covered = [np.zeros(c, dtype=np.int32) for c in counts]
    for unit, frames, first in work:
        if frames == 0:
            continue
        assert 0 <= unit < len(counts) and frames > 0 and first >= 0 and first + frames <= counts[unit]
        covered[unit][first:first + frames] += 1
        if N == 1024 and first > 0:
            assert (firsts[unit] + first) % 2 == 0, "a 1024-point unit is cut in front of an even global frame"
    for c in covered:
        assert (c == 1).all()
    # rounds per group: the busiest one stays close to the mean
    rounds = np.zeros(stride, dtype=np.int64)
    for k, (unit, frames, first) in enumerate(work):
        if frames:
            rounds[k % stride] += -(-(frames + lead) // fpr)
    ideal = sum(-(-(c + lead) // fpr) for c in counts if c > 0)
    per_group = -(-ideal // (sms * groups))
    # every wave cuts a unit per group, and a cut costs the lead-in frames again
    waves = max(1, round(per_group / float(wave))) if wave else 1
    slack = 2 + (waves + 1) * (-(-lead // fpr) + 1)
    assert rounds.max() <= max(per_group, lead // fpr + 1) + slack, (rounds.max(), per_group)
    return rounds


@pytest.mark.parametrize("N,fpr", [(2048, 16), (1024, 32), (4096, 8)])
def test_headline_batch_is_balanced(emul, N, fpr):
    counts = [3000] * 1024
    work, grid = work_list(emul, N, fpr, 1, 148, 1, counts)
    rounds = check(work, grid, N, fpr, 1, 148, 1, counts)
    assert grid == 148
    mean = rounds.sum() / 148.
    assert rounds.max() <= 1.005 * mean + 1, (rounds.max(), mean)      # equal chunks dealt in turn: 1.2 % at best


def test_mixed_lengths_and_few_tracks(emul):
    rng = np.random.default_rng(5)
    for counts in (rng.integers(500, 6000, 1024), rng.integers(500, 6000, 256), [3000] * 64, [3000] * 200,
                   (3000 * rng.lognormal(0, 0.7, 512)).astype(np.int64) + 16):
        work, grid = work_list(emul, 2048, 16, 1, 148, 1, counts)
        rounds = check(work, grid, 2048, 16, 1, 148, 1, counts)
        mean = rounds.sum() / float(len(rounds))
        assert rounds.max() <= 1.02 * mean + 2, (rounds.max(), mean)    # equal chunks: 5 .. 18 % above the mean


def test_small_launches_use_fewer_ctas(emul):
    # one short file: a round per CTA, not a whole chunk on one SM
    work, grid = work_list(emul, 2048, 16, 1, 148, 1, [100])
    rounds = check(work, grid, 2048, 16, 1, 148, 1, [100])
    assert grid == 7 and rounds.max() == 1                              # 100 frames = 15 per round-long piece
    work, grid = work_list(emul, 2048, 16, 1, 148, 1, [1])
    assert grid == 1 and work.tolist() == [[0, 1, 0]]
    work, grid = work_list(emul, 2048, 16, 1, 148, 1, [0, 0])
    assert len(work) == 0
    # 30 s at 100 fps: two rounds per CTA on (nearly) all SMs
    work, grid = work_list(emul, 2048, 16, 1, 148, 1, [3000])
    rounds = check(work, grid, 2048, 16, 1, 148, 1, [3000])
    assert grid >= 90 and rounds.max() == 2


def test_holes_and_empty_units(emul):
    counts = [0, 5, 0, 0, 700, 1, 0, 33, 0]
    work, grid = work_list(emul, 2048, 16, 4, 6, 3, counts)
    check(work, grid, 2048, 16, 4, 6, 3, counts)
    assert not np.isin(work[work[:, 1] > 0, 0], [0, 2, 3, 6, 8]).any()


def test_fixed_chunks_in_unit_order(emul):
    counts = [100, 0, 37, 13]
    work, grid = work_list(emul, 2048, 16, 1, 148, 1, counts, fixed=13)
    assert grid == len(work) == 8 + 3 + 1
    assert work[:8].tolist() == [[0, min(13, 100 - a), a] for a in range(0, 100, 13)]
    assert work[8:].tolist() == [[2, 13, 0], [2, 13, 13], [2, 11, 26], [3, 13, 0]]


@pytest.mark.parametrize("seed", range(60))
def test_random_work_lists(emul, seed):
    rng = np.random.default_rng(1000 + seed)
    N, fpr = [(1024, 32), (1024, 2), (2048, 16), (2048, 1), (4096, 8), (4096, 1)][seed % 6]
    groups = int(rng.choice([1, 2, 4, 16]))
    sms = int(rng.choice([1, 3, 148]))
    lead = int(rng.choice([0, 1, 2, 5, 40]))
    n = int(rng.integers(1, 40))
    scale = int(rng.choice([3, 60, 4000]))
    counts = rng.integers(0, scale, n)
    firsts = rng.integers(0, 1000, n)
    work, grid = work_list(emul, N, fpr, groups, sms, lead, counts, firsts)
    check(work, grid, N, fpr, groups, sms, lead, counts, firsts)


def test_ten_hours_in_one_unit(emul):
    frames = 3_600 * 10 * 100
    # one wave: one run per group
    work, grid = work_list(emul, 2048, 16, 1, 148, 1, [frames], wave=0)
    assert grid == 148 and len(work) == 148
    assert work[:, 1].sum() == frames and (np.diff(work[:, 2]) == work[:-1, 1]).all()
    assert work[:-1, 1].max() - work[:-1, 1].min() == 0
    # waves of 64 rounds: the groups work side by side on a window that moves through the signal
    work, grid = work_list(emul, 2048, 16, 1, 148, 1, [frames])
    rounds = check(work, grid, 2048, 16, 1, 148, 1, [frames])
    assert grid == 148 and work[:, 1].sum() == frames and (np.diff(work[:, 2]) == work[:-1, 1]).all()
    assert rounds[:-1].max() - rounds[:-1].min() <= 1 and rounds[-1] <= rounds[0]


@pytest.mark.parametrize("N,fpr,groups,lead", [(1024, 2, 16, 16), (4096, 1, 8, 16), (2048, 1, 16, 16), (1024, 2, 1, 1)])
def test_lead_in_longer_than_a_round(emul, N, fpr, groups, lead):
    """Single-warp groups: a round holds one or two frames and the lead-in of a cut is many rounds."""
    counts = [200_000, 3, 77_777]
    firsts = [1, 0, 12345]
    work, grid = work_list(emul, N, fpr, groups, 148, lead, counts, firsts)
    check(work, grid, N, fpr, groups, 148, lead, counts, firsts)
    work, grid = work_list(emul, N, fpr, groups, 148, lead, counts, firsts, wave=0)
    check(work, grid, N, fpr, groups, 148, lead, counts, firsts, wave=0)
