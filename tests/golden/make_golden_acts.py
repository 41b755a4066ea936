#!/usr/bin/env python
"""Golden vectors for the peak-picking row beyond ref_sample.npz, computed by the REAL reference
(CPJKU/madmom at /root/reference) through `oracle/ref_loader.py`:

    python tests/golden/make_golden_acts.py      ->  tests/golden/ref_acts.npz

  * the reference's activation fixture tests/data/activations/sample.onsets_rnn.npz (281 values, fps 100), the one
    its tests feed to `peak_picking(act, threshold=0.23, post_max=0)` (tests/test_features_onsets.py:247-253) and to
    the online OnsetPeakPickingProcessor (:267-292);
  * what the reference's `peak_picking` and `OnsetPeakPickingProcessor` return for it and for the SuperFlux fixture
    (offline, with delay, and the online processor fed everything at once -- kept for the record: online mode is
    out of scope on the GPU path).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.ref_loader import load_reference, REFERENCE_ROOT  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
ref = load_reference()
ON = ref.onsets
ACT = os.path.join(REFERENCE_ROOT, "tests", "data", "activations")


def main():
    g = {}
    rnn = np.load(os.path.join(ACT, "sample.onsets_rnn.npz"))
    sf = np.load(os.path.join(ACT, "sample.super_flux.npz"))
    g["rnn_act"], g["rnn_fps"] = np.asarray(rnn["activations"]), np.asarray(rnn["fps"])
    g["sf_act"], g["sf_fps"] = np.asarray(sf["activations"]), np.asarray(sf["fps"])
    rnn_act, sf_act = g["rnn_act"], g["sf_act"]
    rnn_fps, sf_fps = float(g["rnn_fps"]), float(g["sf_fps"])
    g["rnn_peaks_online_params"] = np.asarray(ON.peak_picking(rnn_act, threshold=0.23, post_max=0))
    kw = dict(threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0, combine=0.03)
    g["sf_onsets"] = np.asarray(ON.OnsetPeakPickingProcessor(delay=0, fps=sf_fps, **kw)(sf_act))
    g["sf_onsets_delay1"] = np.asarray(ON.OnsetPeakPickingProcessor(delay=1, fps=sf_fps, **kw)(sf_act))
    g["rnn_onsets_online"] = np.asarray(ON.OnsetPeakPickingProcessor(threshold=0.23, online=True, fps=rnn_fps)(rnn_act))
    # the same parameters the online processor ends up with (smooth = post_avg = post_max = 0), offline
    off = ON.OnsetPeakPickingProcessor(threshold=0.23, smooth=0, post_avg=0, post_max=0, fps=rnn_fps)
    g["rnn_onsets_offline_same_params"] = np.asarray(off(rnn_act))
    # a sweep of combine / delay on the RNN activations
    for combine in (0.0, 0.03, 0.1):
        p = ON.OnsetPeakPickingProcessor(threshold=0.02, smooth=0.02, pre_avg=0.1, post_avg=0.05, pre_max=0.02,
                                         post_max=0.03, combine=combine, delay=0.005, fps=rnn_fps)
        g["rnn_onsets_combine_%g" % combine] = np.asarray(p(rnn_act))
    np.savez_compressed(os.path.join(OUT, "ref_acts.npz"), **g)
    for k, v in g.items():
        print("%-34s %s %s" % (k, v.shape, np.round(v[:8], 4) if v.ndim else v))


if __name__ == "__main__":
    main()
