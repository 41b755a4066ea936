#!/usr/bin/env python
"""Golden vectors for Phase / LocalGroupDelay (madmom/audio/stft.py:136-179, 556-708), computed by the REAL
reference through oracle/ref_loader.py:   python tests/golden/make_golden_phase.py  ->  tests/golden/ref_phase.npz"""
import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.ref_loader import load_reference, REFERENCE_ROOT  # noqa: E402

ref = load_reference()
S, ST = ref.signal, ref.stft
sample_file = os.path.join(REFERENCE_ROOT, "tests", "data", "audio", "sample.wav")
ROWS = [0, 1, 2, 10, 140, 279, 280]
g = {"rows": np.array(ROWS)}
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    ph_plain = ST.Phase(ST.ShortTimeFourierTransform(sample_file))            # the docstring's example (no shift)
g["phase_plain_rows"] = np.asarray(ph_plain)[ROWS]
ph = ST.Phase(sample_file)                                                      # circular_shift defaults to True
assert ph.stft.circular_shift
g["phase_rows"] = np.asarray(ph)[ROWS]
g["lgd_rows"] = np.asarray(ph.lgd())[ROWS]
g["stft_rows"] = np.asarray(ph.stft)[ROWS]
g["lgd_from_class_rows"] = np.asarray(ST.LocalGroupDelay(sample_file))[ROWS]
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_phase.npz"), **g)
print("ref_phase.npz", {k: v.shape for k, v in g.items()}, g["phase_rows"].dtype, g["lgd_rows"].dtype)
