#!/usr/bin/env python
"""Golden vectors for SURVEY 8(f) rows 2 and 3, computed by the REAL reference (CPJKU/madmom at /root/reference)
through `oracle/ref_loader.py`:

    python tests/golden/make_golden_mel.py      ->  tests/golden/ref_mel.npz

  * MelFilterbank band tables (madmom/audio/filters.py:1035-1092): the 40-band default on 1024 bins and the
    CNNOnsetProcessor bank (80 bands, fmin 27.5, fmax 16000, norm, unique_filters=False) for N = 1024/2048/4096;
    BarkFilterbank ('normal') dense matrix as a non-triangular case;
  * the reference's FilteredSpectrogram(MelFilterbank, 40) on sample.wav (tests/test_audio_spectrogram.py:173-188);
  * the CNN onset front end of madmom/features/onsets.py:838-871 on sample.wav: np.dstack over frame sizes
    [2048, 1024, 4096] of log(filt + EPSILON), (281, 80, 3);
  * gain / attenuation / normalisation of the ingest path (madmom/audio/signal.py:71-167) on int16 and float32.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.ref_loader import load_reference, REFERENCE_ROOT  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
ref = load_reference()
S, ST, FI, SP, P = ref.signal, ref.stft, ref.filters, ref.spectrogram, ref.processors
sample_file = os.path.join(REFERENCE_ROOT, "tests", "data", "audio", "sample.wav")
EPSILON = np.spacing(1)


def band_table(fb):
    fb = np.asarray(fb)
    starts, lens, weights = [], [], []
    for b in range(fb.shape[1]):
        nz = np.nonzero(fb[:, b])[0]
        assert len(nz) and np.all(np.diff(nz) == 1)
        starts.append(nz[0]), lens.append(len(nz)), weights.append(fb[nz, b])
    return (np.array(starts, np.int32), np.array(lens, np.int32), np.concatenate(weights).astype(np.float32))


def main():
    g = {}
    fb = FI.MelFilterbank(ST.fft_frequencies(1024, 44100))                   # 40 bands, 20 .. 17000 Hz
    g["mel40_N2048_start"], g["mel40_N2048_len"], g["mel40_N2048_w"] = band_table(fb)
    g["mel40_N2048_center"] = np.asarray(fb.center_frequencies)
    for N in (1024, 2048, 4096):
        fb = FI.MelFilterbank(ST.fft_frequencies(N // 2, 44100), num_bands=80, fmin=27.5, fmax=16000,
                              norm_filters=True, unique_filters=False)
        g["mel80_N%d_dense" % N] = np.asarray(fb)                            # duplicate low bands: keep it dense
        g["mel80_N%d_center" % N] = np.asarray(fb.center_frequencies)
    # the bank FilteredSpectrogram(filterbank=BarkFilterbank, num_bands='normal') builds (its own fmin / fmax defaults)
    g["bark_N2048_dense"] = np.asarray(FI.BarkFilterbank(ST.fft_frequencies(1024, 44100), num_bands="normal",
                                                         fmin=30., fmax=17000., norm_filters=True,
                                                         unique_filters=True))
    sig = S.Signal(sample_file)
    g["filt_mel40"] = np.asarray(SP.FilteredSpectrogram(sample_file, filterbank=FI.MelFilterbank, num_bands=40))
    g["filt_bark"] = np.asarray(SP.FilteredSpectrogram(sample_file, filterbank=FI.BarkFilterbank,
                                                       num_bands="normal"))
    multi = P.ParallelProcessor([])
    for frame_size in [2048, 1024, 4096]:
        frames = S.FramedSignalProcessor(frame_size=frame_size, fps=100)
        stft = ST.ShortTimeFourierTransformProcessor()
        filt = SP.FilteredSpectrogramProcessor(filterbank=FI.MelFilterbank, num_bands=80, fmin=27.5, fmax=16000,
                                               norm_filters=True, unique_filters=False)
        spec = SP.LogarithmicSpectrogramProcessor(log=np.log, add=EPSILON)
        multi.append(P.SequentialProcessor((frames, stft, filt, spec)))
    pre = P.SequentialProcessor((S.SignalProcessor(num_channels=1, sample_rate=44100), multi, np.dstack))
    g["cnn_onset_frontend"] = np.asarray(pre(sample_file)).astype(np.float32)
    assert g["cnn_onset_frontend"].shape == (281, 80, 3)
    # ---- ingest: gain / attenuation / normalisation (signal.py:71-167)
    x16 = np.asarray(sig)[:5000].copy()
    xf = (x16 / 32768.).astype(np.float32)
    for tag, x in (("i16", x16), ("f32", xf)):
        for gain in (-6.5, 0.0, 3.0):
            if gain > 0 and tag == "i16":
                continue                                                     # the reference raises ValueError
            g["gain_%s_%g" % (tag, gain)] = np.asarray(S.adjust_gain(x, gain))
        g["atten_%s_10" % tag] = np.asarray(S.attenuate(x, 10.))
    g["norm_f32"] = np.asarray(S.normalize(xf))
    g["norm_i16"] = np.asarray(S.normalize(x16))
    g["ingest_i16"], g["ingest_f32"] = x16, xf
    np.savez_compressed(os.path.join(OUT, "ref_mel.npz"), **g)
    print("ref_mel.npz %.1f KB" % (os.path.getsize(os.path.join(OUT, "ref_mel.npz")) / 1024.))


if __name__ == "__main__":
    main()
