"""bench.py --impl reference (CPU only): the arm the driver times beside the GPU arm.  It must drive the reference's
own classes (copy under baseline/_ref made by __graft_entry__.build(), or /root/reference through the same loader),
print ONE JSON line with the contract's keys and the same `config` keys as the GPU arm, and do nothing on ranks
other than 0."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, *args):
    env = dict(os.environ)
    env.update(extra_env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--seconds", "1",
                           "--steps", "1", "--warmup", "0"] + list(args), capture_output=True, text=True, env=env,
                          timeout=600, cwd=ROOT)


def test_reference_arm_line():
    res = _run()
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["unit"] == "audio-s/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["gpu_launches"] == 0
    assert line["e2e"] == {"value": line["value"], "unit": "audio-s/s", "h2d_bytes_per_step": 0,
                           "d2h_bytes_per_step": 0}
    cb = line["cpu_baseline"]
    assert cb["value"] == line["value"] and cb["cores"] >= 1 and cb["sample"]
    have_copy = os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "madmom", "audio"))
    assert cb["kind"] == ("reference" if have_copy else "port")
    # the keys the driver compares between the two arms
    for key in ("workload", "tracks_per_gpu", "seconds", "pcm", "sample_rate"):
        assert key in line["config"], key
    assert line["config"]["workload"] == "logfiltspecflux"


def test_reference_arm_is_silent_on_other_ranks():
    res = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert res.returncode == 0 and res.stdout.strip() == ""


@pytest.mark.parametrize("workload", ["logfiltspecflux", "superflux", "multires_beat", "complexflux"])
def test_reference_processors_agree_with_the_oracle(workload):
    """What the reference arm times is the same computation the parity tests check: the reference's own chain on
    a seeded track equals the oracle's result for that workload."""
    sys.path.insert(0, ROOT)
    import bench
    if not bench.reference_copy_available():
        pytest.skip("baseline/_ref not built (run __graft_entry__.build() where /root/reference exists)")
    from madmom_b200.synth import make_track
    from oracle import frontend as O
    x = make_track(77, 1.5)
    got = np.asarray(bench._reference_processor(workload)(x))
    if workload == "logfiltspecflux":
        want = O.run_recipe(x, O.LOGFILTSPECFLUX, stages=True)["odf"]
    elif workload == "superflux":
        want = O.run_recipe(x, O.SUPERFLUX, stages=True)["odf"]
    elif workload == "complexflux":
        st = O.run_recipe(x, dict(O.SUPERFLUX_200, circular_shift=True), stages=True)
        want = O.complex_flux(st["log"], st["stft"], st["filterbank"], st["diff_frames"])
    else:
        cols = []
        for N in (1024, 2048, 4096):
            st = O.run_recipe(x, dict(frame_size=N, hop_size=441., num_bands=6, norm_filters=True, mul=1, add=1,
                                      diff_ratio=0.5, positive_diffs=True), stages=True)
            cols += [st["log"], st["diff"]]
        want = np.hstack(cols)
    assert got.shape == want.shape
    assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
