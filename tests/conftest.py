import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


def _npz(name):
    with np.load(os.path.join(GOLDEN, name)) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def sample_pcm():
    return _npz("sample_wav.npz")["pcm"]


@pytest.fixture(scope="session")
def ref_sample():
    return _npz("ref_sample.npz")


@pytest.fixture(scope="session")
def ref_filterbanks():
    return _npz("ref_filterbanks.npz")


@pytest.fixture(scope="session")
def ref_framing():
    return _npz("ref_framing.npz")


@pytest.fixture(scope="session")
def ref_synth():
    return _npz("ref_synth.npz")


@pytest.fixture(scope="session")
def ref_mel():
    return _npz("ref_mel.npz")


@pytest.fixture(scope="session")
def ref_phase():
    return _npz("ref_phase.npz")


@pytest.fixture(scope="session")
def ref_acts():
    return _npz("ref_acts.npz")


@pytest.fixture(scope="session")
def ref_odfs():
    return _npz("ref_odfs.npz")


def assert_close(actual, desired, rtol=1e-4, atol=1e-5, what=""):
    """The parity bound of BASELINE.md section 2: |a - b| <= atol + rtol * |b| element-wise."""
    actual = np.asarray(actual)
    desired = np.asarray(desired)
    assert actual.shape == desired.shape, "%s shape %s != %s" % (what, actual.shape, desired.shape)
    err = np.abs(actual.astype(np.float64) - desired.astype(np.float64))
    bound = atol + rtol * np.abs(desired.astype(np.float64))
    bad = err > bound
    if bad.any():
        i = np.unravel_index(np.argmax(err - bound), err.shape)
        raise AssertionError("%s: %d/%d elements out of tolerance; worst at %s: got %r want %r (err %.3g > %.3g)"
                             % (what, bad.sum(), bad.size, i, actual[i], desired[i], err[i], bound[i]))
