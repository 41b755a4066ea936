#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running the REAL reference (CPJKU/madmom, mounted
read-only at /root/reference) through `oracle/ref_loader.py`.

Run from the repo root in the build container (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

Outputs (all small, committed):
    sample_wav.npz      the reference's tests/data/audio/sample.wav as an int16 array (input fixture)
    ref_sample.npz      every stage of the path on sample.wav computed by the reference, plus the
                        reference's shipped fixtures (tests/data/activations/*.npz, detections/*.txt)
    ref_filterbanks.npz LogarithmicFilterbank for N in {1024,2048,4096} x {6 norm, 12, 24} bands/octave
    ref_framing.npz     num_frames / frame-start / diff_frames tables (integer, bit-exact)
    ref_synth.npz       the recipes of BASELINE.json on seeded synthetic tracks (madmom_b200.synth)
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
S, ST, FI, SP, ON = ref.signal, ref.stft, ref.filters, ref.spectrogram, ref.onsets
DATA = os.path.join(REFERENCE_ROOT, "tests", "data")
sample_file = os.path.join(DATA, "audio", "sample.wav")

ROWS = [0, 1, 2, 10, 140, 279, 280]


def sig_from(array, sample_rate=44100):
    return S.Signal(array, sample_rate=sample_rate, num_channels=1)


def multires(sig, mul, diff_ratio, fps=100):
    """The RNNOnsetProcessor / RNNBeatProcessor pre-processor (features/onsets.py:767-791,
    features/beats.py:80-106) built from the reference's own processors."""
    P = ref.processors
    multi = P.ParallelProcessor([])
    for frame_size in [1024, 2048, 4096]:
        frames = S.FramedSignalProcessor(frame_size=frame_size, fps=fps)
        stft = ST.ShortTimeFourierTransformProcessor()
        filt = SP.FilteredSpectrogramProcessor(num_bands=6, fmin=30, fmax=17000, norm_filters=True)
        spec = SP.LogarithmicSpectrogramProcessor(mul=mul, add=1)
        diff = SP.SpectrogramDifferenceProcessor(diff_ratio=diff_ratio, positive_diffs=True,
                                                 stack_diffs=np.hstack)
        multi.append(P.SequentialProcessor((frames, stft, filt, spec, diff)))
    return np.asarray(P.SequentialProcessor((multi, np.hstack))(sig))


def main():
    # ---------------------------------------------------------------- sample.wav input fixture
    sig = S.Signal(sample_file)
    np.savez_compressed(os.path.join(OUT, "sample_wav.npz"), pcm=np.asarray(sig),
                        sample_rate=np.int64(sig.sample_rate))

    # ---------------------------------------------------------------- stages on sample.wav
    g = {}
    frames = S.FramedSignal(sig, frame_size=2048, hop_size=441.)
    g["num_frames_100"] = np.int64(frames.num_frames)
    frames200 = S.FramedSignal(sig, frame_size=2048, fps=200)
    g["num_frames_200"] = np.int64(frames200.num_frames)
    g["rows"] = np.array(ROWS)
    g["frames_rows"] = np.array([np.asarray(frames[i]) for i in ROWS])
    stft = ST.ShortTimeFourierTransform(frames)
    g["stft_rows"] = np.asarray(stft)[ROWS]
    g["fft_window"] = np.asarray(stft.fft_window)
    stft_cs = ST.ShortTimeFourierTransform(frames, circular_shift=True)
    g["stft_circ_rows"] = np.asarray(stft_cs)[ROWS]
    spec = SP.Spectrogram(stft)
    g["spec_rows"] = np.asarray(spec)[ROWS]
    g["spec_col_sums"] = np.asarray(spec, dtype=np.float64).sum(axis=0)
    filt = SP.FilteredSpectrogram(spec)                       # 12 bands, norm_filters=True (defaults)
    g["filt12n"] = np.asarray(filt)
    g["logfilt12n"] = np.asarray(SP.LogarithmicFilteredSpectrogram(spec))
    g["logspec_rows"] = np.asarray(SP.LogarithmicSpectrogram(spec))[ROWS]
    # SuperFlux (config 1): 24 bands, no norm, max filter 3, d=1
    sf = SP.SuperFluxProcessor()(frames)
    g["superflux_diff"] = np.asarray(sf)
    lf24 = SP.LogarithmicFilteredSpectrogram(spec, num_bands=24, norm_filters=False)
    g["logfilt24"] = np.asarray(lf24)
    g["superflux_keepdims"] = np.asarray(lf24.diff(diff_max_bins=3, positive_diffs=True))
    g["odf_superflux_100"] = np.asarray(ON.superflux(lf24))
    # signed (non-positive) difference with max filter, explicit diff_frames=2
    g["diff_signed_d2"] = np.asarray(lf24.diff(diff_frames=2, diff_max_bins=3, positive_diffs=False))
    # LogFiltSpecFlux program recipe (config 2): 12 bands, no norm, d=1, spectral_flux
    lf12 = SP.LogarithmicFilteredSpectrogram(spec, num_bands=12, norm_filters=False)
    g["logfilt12"] = np.asarray(lf12)
    g["odf_logfiltspecflux"] = np.asarray(ON.spectral_flux(lf12))
    g["odf_spectralflux_plain"] = np.asarray(ON.spectral_flux(spec))
    # programs at fps 200 (bin/SuperFlux, bin/ComplexFlux)
    kw = dict(fps=200, filterbank=FI.LogarithmicFilterbank, num_bands=24, fmin=30, fmax=17000,
              norm_filters=False, log=np.log10, mul=1, add=1, diff_ratio=0.5, diff_max_bins=3,
              positive_diffs=True)
    g["odf_superflux_200"] = np.asarray(ON.SpectralOnsetProcessor(onset_method="superflux", **kw)(sample_file))
    g["odf_complexflux_200"] = np.asarray(
        ON.SpectralOnsetProcessor(onset_method="complex_flux", **kw)(sample_file))
    spec200c = SP.Spectrogram(ST.ShortTimeFourierTransform(frames200, circular_shift=True))
    lgd = np.asarray(spec200c.stft.phase().lgd())
    g["lgd_200_rows"] = lgd[ROWS]
    # shipped fixtures of the reference
    for name in ("super_flux", "log_filt_spec_flux", "complex_flux", "spectral_flux"):
        act = np.load(os.path.join(DATA, "activations", "sample.%s.npz" % name))
        g["fixture_act_" + name] = act["activations"]
        g["fixture_fps_" + name] = act["fps"]
    for name in ("super_flux", "log_filt_spec_flux", "complex_flux"):
        g["fixture_det_" + name] = np.loadtxt(os.path.join(DATA, "detections", "sample.%s.txt" % name))
    # peak picking through the reference's OnsetPeakPickingProcessor on the reference ODFs
    pp = ON.OnsetPeakPickingProcessor(threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                                      combine=0.03, delay=0, fps=200)
    g["onsets_superflux_200"] = np.asarray(pp(g["odf_superflux_200"]))
    g["peaks_superflux_200"] = ON.peak_picking(g["odf_superflux_200"], 1.1, None, 30, 0, 2, 10)
    pp = ON.OnsetPeakPickingProcessor(threshold=1.6, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                                      combine=0.03, delay=0, fps=100)
    g["onsets_logfiltspecflux"] = np.asarray(pp(g["odf_logfiltspecflux"]))
    pp = ON.OnsetPeakPickingProcessor(threshold=0.25, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                                      combine=0.03, delay=0, fps=200)
    g["onsets_complexflux_200"] = np.asarray(pp(g["odf_complexflux_200"]))
    # multi-resolution stacks (config 3)
    g["rnn_onset_stack"] = multires(sig_from(np.asarray(sig)), mul=5, diff_ratio=0.25)
    g["rnn_beat_stack"] = multires(sig_from(np.asarray(sig)), mul=1, diff_ratio=0.5)
    np.savez_compressed(os.path.join(OUT, "ref_sample.npz"), **g)

    # ---------------------------------------------------------------- filterbanks
    fbs = {}
    for N in (1024, 2048, 4096):
        binf = ST.fft_frequencies(N // 2, 44100)
        for bpo, norm in ((6, True), (12, False), (12, True), (24, False)):
            fb = np.asarray(FI.LogarithmicFilterbank(binf, num_bands=bpo, fmin=30, fmax=17000,
                                                     norm_filters=norm, unique_filters=True))
            key = "N%d_b%d_%s" % (N, bpo, "norm" if norm else "raw")
            starts, lens, weights = [], [], []
            for b in range(fb.shape[1]):
                nz = np.nonzero(fb[:, b])[0]
                assert np.all(np.diff(nz) == 1)
                starts.append(nz[0])
                lens.append(len(nz))
                weights.append(fb[nz, b])
            fbs[key + "_start"] = np.array(starts, dtype=np.int32)
            fbs[key + "_len"] = np.array(lens, dtype=np.int32)
            fbs[key + "_w"] = np.concatenate(weights).astype(np.float32)
            fbs[key + "_center"] = np.asarray(
                FI.LogarithmicFilterbank(binf, num_bands=bpo, fmin=30, fmax=17000, norm_filters=norm,
                                         unique_filters=True).center_frequencies)
    # non-unique filters variant (duplicate low bands)
    fb = np.asarray(FI.LogarithmicFilterbank(ST.fft_frequencies(1024, 44100), num_bands=12,
                                             unique_filters=False))
    fbs["N2048_b12_norm_nonunique_dense"] = fb
    np.savez_compressed(os.path.join(OUT, "ref_filterbanks.npz"), **fbs)

    # ---------------------------------------------------------------- integer framing tables
    fr = {}
    cases = []
    rng = np.random.default_rng(7)
    lens = [1, 2, 440, 441, 442, 882, 1000, 123481, 1323000, 1587600000] + list(rng.integers(1, 5_000_000, 20))
    for n in lens:
        for hop in (441., 220.5, 512., 100.25, 160.):
            for end in ("normal", "extend"):
                # zero-stride view: a signal of length n that costs no memory
                fake = np.lib.stride_tricks.as_strided(np.zeros(1, np.int16), shape=(int(n),), strides=(0,))
                fsig = S.FramedSignal(S.Signal(fake, sample_rate=44100), frame_size=2048, hop_size=hop,
                                      end=end)
                nf = fsig.num_frames
                cases.append((int(n), hop, 0 if end == "normal" else 1, nf))
    fr["num_frames_cases"] = np.array(cases, dtype=np.float64)
    # frame contents via the reference's signal_frame on a ramp signal: recover starts + padding exactly
    ramp = (np.arange(20000) % 30000 + 1).astype(np.int16)
    st_cases = []
    for N in (1024, 2048, 4096):
        for hop in (441., 220.5, 100.25):
            for origin in (0, "online", "stream", 100, -37):
                fs = S.FramedSignal(sig_from(ramp), frame_size=N, hop_size=hop, origin=origin)
                for idx in list(range(0, 12)) + list(range(fs.num_frames - 6, fs.num_frames)):
                    frame = np.asarray(fs[idx])
                    nzp = np.nonzero(frame)[0]
                    first_nz = int(nzp[0]) if len(nzp) else -1
                    first_val = int(frame[nzp[0]]) if len(nzp) else 0
                    st_cases.append((N, hop, fs.origin, idx, fs.num_frames, first_nz, first_val,
                                     int(len(nzp)), int(frame.astype(np.int64).sum())))
    fr["frame_cases"] = np.array(st_cases, dtype=np.float64)
    df = []
    for N in (512, 1024, 2048, 4096, 8192):
        for hop in (441., 220.5, 882.):
            for ratio in (0.25, 0.5, 0.75):
                df.append((N, hop, ratio, SP._diff_frames(ratio, hop, N, np.hanning)))
    fr["diff_frames_cases"] = np.array(df, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "ref_framing.npz"), **fr)

    # ---------------------------------------------------------------- synthetic tracks
    sy = {}
    for seed, seconds in ((1000, 3.0), (1001, 2.5)):
        x = make_track(seed, seconds)
        sg = sig_from(x)
        tag = "s%d" % seed
        # C1 SuperFlux @100 fps
        fs = S.FramedSignalProcessor(frame_size=2048, fps=100)(sg)
        sy[tag + "_superflux"] = np.asarray(SP.SuperFluxProcessor()(fs))
        # C2 LogFiltSpecFlux: 12 bands no norm, d=1, positive diffs
        sp = SP.Spectrogram(fs)
        lf = SP.LogarithmicFilteredSpectrogram(sp, num_bands=12, norm_filters=False)
        dd = lf.diff(positive_diffs=True)
        sy[tag + "_lfsf_log"] = np.asarray(lf)
        sy[tag + "_lfsf_diff"] = np.asarray(dd)
        sy[tag + "_lfsf_odf"] = np.asarray(ON.spectral_flux(lf))
        # SuperFlux program @200 fps
        fs2 = S.FramedSignalProcessor(frame_size=2048, fps=200)(sg)
        lf2 = SP.LogarithmicFilteredSpectrogram(SP.Spectrogram(fs2), num_bands=24, norm_filters=False)
        sy[tag + "_superflux200_odf"] = np.asarray(ON.superflux(lf2))
        # C3 multi-resolution stacks
        sy[tag + "_rnn_onset"] = multires(sg, mul=5, diff_ratio=0.25)
        sy[tag + "_rnn_beat"] = multires(sg, mul=1, diff_ratio=0.5)
    # float32 input variant of seed 1000 (window unscaled)
    xf = make_track(1000, 3.0, dtype=np.float32)
    fs = S.FramedSignalProcessor(frame_size=2048, fps=100)(sig_from(xf))
    sy["s1000_f32_superflux"] = np.asarray(SP.SuperFluxProcessor()(fs))
    np.savez_compressed(os.path.join(OUT, "ref_synth.npz"), **sy)
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print("%-24s %8.1f KB" % (f, os.path.getsize(os.path.join(OUT, f)) / 1024.))


if __name__ == "__main__":
    main()
