#!/usr/bin/env python
"""Golden vectors for the spectral onset detection functions of SURVEY.md 8(f) row 4, computed by the REAL
reference (CPJKU/madmom, read-only at /root/reference) through `oracle/ref_loader.py`.

    python tests/golden/make_golden_odfs.py          # build container only (the GPU box has no /root/reference)

Output: tests/golden/ref_odfs.npz -- every function of madmom/features/onsets.py:102-160,320-581 on the
reference's tests/data/audio/sample.wav with the inputs its own tests use (tests/test_features_onsets.py:24-27):
    sample_spec          = Spectrogram(sample.wav, circular_shift=True)
    sample_log_filt_spec = LogarithmicFilteredSpectrogram(sample_spec, num_bands=24, mul=1, add=1)
and the same functions on a seeded synthetic track (madmom_b200.synth.make_track(3, 2.0)).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.ref_loader import load_reference, REFERENCE_ROOT  # noqa: E402
from madmom_b200.synth import make_track  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
ref = load_reference()
S, SP, ON = ref.signal, ref.spectrogram, ref.onsets
sample_file = os.path.join(REFERENCE_ROOT, "tests", "data", "audio", "sample.wav")


def odfs(prefix, spec, log_filt_spec, out):
    out[prefix + "hfc_logfilt"] = ON.high_frequency_content(log_filt_spec)
    out[prefix + "hfc_spec"] = ON.high_frequency_content(spec)
    out[prefix + "sd_logfilt"] = ON.spectral_diff(log_filt_spec)
    out[prefix + "sd_spec"] = ON.spectral_diff(spec)
    out[prefix + "mkl_logfilt"] = ON.modified_kullback_leibler(log_filt_spec)
    out[prefix + "mkl_spec"] = ON.modified_kullback_leibler(spec)
    out[prefix + "pd"] = ON.phase_deviation(log_filt_spec)
    out[prefix + "wpd"] = ON.weighted_phase_deviation(spec)
    out[prefix + "nwpd"] = ON.normalized_weighted_phase_deviation(spec)
    out[prefix + "cd"] = ON.complex_domain(spec)
    out[prefix + "rcd"] = ON.rectified_complex_domain(spec)


def main():
    out = {}
    spec = SP.Spectrogram(sample_file, circular_shift=True)
    lfs = SP.LogarithmicFilteredSpectrogram(spec, num_bands=24, mul=1, add=1)
    odfs("sample_", spec, lfs, out)
    x = make_track(3, 2.0)
    sig = S.Signal(x, sample_rate=44100, num_channels=1)
    spec = SP.Spectrogram(sig, circular_shift=True)
    lfs = SP.LogarithmicFilteredSpectrogram(spec, num_bands=24, mul=1, add=1)
    odfs("synth_", spec, lfs, out)
    for k, v in out.items():
        out[k] = np.asarray(v)
        print("%-22s %s %s  %s" % (k, out[k].dtype, out[k].shape, np.array2string(out[k][:4], precision=6)))
    np.savez_compressed(os.path.join(OUT, "ref_odfs.npz"), **out)


if __name__ == "__main__":
    main()
