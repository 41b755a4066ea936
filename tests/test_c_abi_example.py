"""The C ABI used from plain C (tests/c_abi_example.c): compiles and links without Python or CUDA headers (CPU), and
on a GPU prints the same numbers as the ctypes path for the same signal."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_DIR = os.path.join(ROOT, "madmom_b200", "lib")
SRC = os.path.join(ROOT, "tests", "c_abi_example.c")


def _build(tmp_path):
    if not os.path.exists(os.path.join(LIB_DIR, "libmadmom_b200.so")):
        sys.path.insert(0, ROOT)
        from madmom_b200 import build
        build.build_library()
    exe = str(tmp_path / "c_abi_example")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-O1", SRC, "-o", exe, "-L" + LIB_DIR,
           "-lmadmom_b200", "-Wl,-rpath," + LIB_DIR]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return exe


def test_header_is_c99_and_links(tmp_path):
    exe = _build(tmp_path)
    assert os.path.exists(exe)
    # without arguments the program exits before touching the GPU
    assert subprocess.run([exe]).returncode == 2


@pytest.mark.gpu
def test_c_program_matches_python_path(tmp_path):
    torch = pytest.importorskip("torch")
    from madmom_b200 import recipe as R
    from madmom_b200.batch import BatchFrontEnd
    exe = _build(tmp_path)
    fb = R.SUPERFLUX.build_filterbank()
    starts, lens, weights = fb.band_table()
    path = str(tmp_path / "recipe.bin")
    with open(path, "wb") as f:
        np.int32(len(starts)).tofile(f)
        starts.astype(np.int32).tofile(f)
        lens.astype(np.int32).tofile(f)
        np.asarray(weights, dtype=np.float32).tofile(f)
        R.fft_window(np.hanning, 2048, np.int16).astype(np.float64).tofile(f)
    from madmom_b200.synth import make_track
    n = 3 * 44100 + 123
    pcm = make_track(21, num_samples=n)
    raw = str(tmp_path / "pcm.raw")
    pcm.astype("<i2").tofile(raw)
    res = subprocess.run([exe, path, raw], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    head, nums = res.stdout.strip().splitlines()
    tok = head.split()
    assert int(tok[tok.index("frames") + 1]) == R.num_frames(n, 441.) and int(tok[tok.index("launches") + 1]) == 1
    got = {k: float(v) for k, v in zip(nums.split()[0::2], nums.split()[1::2])}
    # the same signal through the Python binding
    out = BatchFrontEnd(R.SUPERFLUX).run_host(pcm, [n], ("diff", "odf"))
    diff, odf = out["diff"].numpy(), out["odf"].numpy()
    want = {"sum_diff": float(diff.astype(np.float64).sum()), "sum_odf": float(odf.astype(np.float64).sum()),
            "odf10": float(odf[10]), "odf_last": float(odf[-1])}
    for k in want:
        assert abs(got[k] - want[k]) <= 1e-8 * max(1.0, abs(want[k])), (k, got[k], want[k])
