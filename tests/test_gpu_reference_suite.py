"""The reference's own test classes for the spectrogram stages, asserted against the drop-in classes: one test per
class of /root/reference/tests/test_audio_spectrogram.py (lines cited per test), with the same type checks, attribute
values, methods, processor defaults and literal spot values (the reference's literals are known answers of the path,
SURVEY 8(c)).  `sig` (the samples of the reference's sample.wav, tests/golden/sample_wav.npz) stands in for the file
path: the file itself does not travel to the GPU box.  Literals are compared with np.allclose defaults, like there."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

SPEC_HEAD = [3.15249, 4.00272, 5.66156, 6.30141, 6.02199, 10.84909, 17.83130, 19.44511]       # :85-87
SPEC_TAIL = [0.0365325, 0.036513, 0.0364213, 0.0366203, 0.036737, 0.036423, 0.036335, 0.0367054]
FILT_HEAD = [5.661564, 6.30141, 6.02199, 10.84909, 17.8313, 19.44511, 17.56456, 21.859523]   # :149-151
FILT_TAIL = [0.123125, 0.119462, 0.137849, 0.1269156, 0.110888, 0.083526, 0.05426, 0.064614]
LOG_HEAD = [0.618309, 0.699206, 0.823576, 0.86341, 0.84646, 1.073685, 1.27488, 1.310589]      # :287-289
LOG_TAIL = [0.015583, 0.0155747, 0.0155363, 0.0156197, 0.0156684, 0.015537, 0.0155003, 0.01565535]
LOGFILT_HEAD = [0.8235762, 0.863407, 0.8464602, 1.073685, 1.27488, 1.3105896, 1.2686847, 1.359067]     # :366-368
LOGFILT_TAIL = [0.05042794, 0.0490095, 0.05608485, 0.05189138, 0.04567042, 0.03483925, 0.02294769, 0.02719229]


@pytest.fixture(scope="module")
def sig(sample_pcm):
    from madmom_b200.audio.signal import Signal
    return Signal(sample_pcm, sample_rate=44100)


@pytest.fixture(scope="module")
def fft_freqs():
    from madmom_b200.audio.stft import fft_frequencies
    return fft_frequencies(1024, 44100)


@pytest.fixture(scope="module")
def log_centres(ref_filterbanks):
    return ref_filterbanks["N2048_b12_norm_center"]             # LOG_FILTERBANK_CENTER_FREQS of the reference's tests


def test_spectrogram_class(sig, fft_freqs):
    """TestSpectrogramClass, :70-107."""
    from madmom_b200.audio.stft import ShortTimeFourierTransform
    from madmom_b200.audio.spectrogram import (Spectrogram, SpectrogramDifference, FilteredSpectrogram,
                                               LogarithmicSpectrogram)
    result = Spectrogram(sig)
    assert isinstance(result, Spectrogram) and isinstance(result.stft, ShortTimeFourierTransform)
    assert isinstance(result.bin_frequencies, np.ndarray)
    assert isinstance(result.num_bins, int) and isinstance(result.num_frames, int)
    assert np.allclose(result[0, :8], SPEC_HEAD) and np.allclose(result[0, -8:], SPEC_TAIL)
    assert result.shape == (281, 1024) and np.allclose(result.bin_frequencies, fft_freqs)
    assert result.num_frames == 281 and result.num_bins == 1024
    assert np.allclose(np.asarray(Spectrogram(result)), np.asarray(result))                     # from spec
    assert np.allclose(np.asarray(Spectrogram(ShortTimeFourierTransform(sig))), np.asarray(result))   # from stft
    assert isinstance(result.diff(), SpectrogramDifference)
    assert isinstance(result.filter(), FilteredSpectrogram)
    assert isinstance(result.log(), LogarithmicSpectrogram)
    from madmom_b200.audio.spectrogram import spec              # the module function keeps a device STFT on the device
    assert isinstance(spec(ShortTimeFourierTransform(sig)), Spectrogram)


def test_spectrogram_processor_class(sig, fft_freqs):
    """TestSpectrogramProcessorClass, :110-126."""
    from madmom_b200.audio.spectrogram import Spectrogram, SpectrogramProcessor
    result = SpectrogramProcessor().process(sig)
    assert isinstance(result, Spectrogram)
    assert result.shape == (281, 1024) and np.allclose(result.bin_frequencies, fft_freqs)
    assert result.num_frames == 281 and result.num_bins == 1024


def test_filtered_spectrogram_class(sig, log_centres):
    """TestFilteredSpectrogramClass, :129-206 (the Mel / Bark values of test_filterbanks are in
    test_gpu_processors.py::test_mel_filterbank_and_cnn_front_end)."""
    from madmom_b200.audio.stft import ShortTimeFourierTransform
    from madmom_b200.audio.filters import LogarithmicFilterbank
    from madmom_b200.audio.spectrogram import (Spectrogram, FilteredSpectrogram, SpectrogramDifference,
                                               LogarithmicSpectrogram)
    result = FilteredSpectrogram(sig)
    assert isinstance(result, FilteredSpectrogram) and isinstance(result, Spectrogram)
    assert isinstance(result.stft, ShortTimeFourierTransform)
    assert isinstance(result.filterbank, LogarithmicFilterbank)
    assert isinstance(result.bin_frequencies, np.ndarray)
    assert isinstance(result.num_bins, int) and isinstance(result.num_frames, int)
    with pytest.raises(TypeError):
        FilteredSpectrogram(sig, filterbank="bla")
    assert np.allclose(result[0, :8], FILT_HEAD) and np.allclose(result[0, -8:], FILT_TAIL)
    assert result.shape == (281, 81) and np.allclose(result.bin_frequencies, log_centres)
    assert result.num_bins == 81 and result.num_frames == 281
    again = FilteredSpectrogram(sig, filterbank=result.filterbank)                     # with given filterbank
    assert again.shape == (281, 81) and np.allclose(again.bin_frequencies, log_centres)
    assert again.num_bins == 81 and again.num_frames == 281
    spec = Spectrogram(sig)                                                            # test_from_spec
    from_spec = FilteredSpectrogram(spec)
    assert np.allclose(from_spec[0, :8], FILT_HEAD)
    assert np.allclose(spec[0, :8], SPEC_HEAD)                                         # spec must not be altered
    assert isinstance(result.diff(), SpectrogramDifference)
    assert isinstance(result.log(), LogarithmicSpectrogram)


def test_filtered_spectrogram_processor_class(sig, log_centres):
    """TestFilteredSpectrogramProcessorClass, :209-267."""
    from madmom_b200.audio.filters import LogarithmicFilterbank
    from madmom_b200.audio.spectrogram import FilteredSpectrogram, FilteredSpectrogramProcessor
    proc = FilteredSpectrogramProcessor()
    assert issubclass(proc.filterbank, LogarithmicFilterbank)
    assert isinstance(proc.num_bands, int) and isinstance(proc.fmin, float) and isinstance(proc.fmax, float)
    assert isinstance(proc.fref, float) and isinstance(proc.norm_filters, bool)
    assert isinstance(proc.unique_filters, bool)
    assert proc.num_bands == 12 and proc.fmin == 30 and proc.fmax == 17000 and proc.fref == 440
    assert proc.norm_filters is True and proc.unique_filters is True
    result = proc.process(sig)
    assert isinstance(result, FilteredSpectrogram)
    assert result.shape == (281, 81) and np.allclose(result.bin_frequencies, log_centres)
    assert result.num_bins == 81 and result.num_frames == 281
    other = FilteredSpectrogramProcessor(num_bands=6, fmin=300, fmax=10000).process(sig)        # test_other_values
    assert isinstance(other, FilteredSpectrogram) and other.shape == (281, 29)
    assert np.allclose(other.bin_frequencies,
                       [344.53125, 387.5976562, 430.6640625, 495.2636718, 559.86328125, 624.4628906, 689.0625,
                        775.1953125, 882.8613281, 990.52734375, 1098.1933593, 1248.9257812, 1399.6582031,
                        1571.9238281, 1765.7226562, 1981.0546875, 2217.9199218, 2497.8515625, 2799.3164062,
                        3143.84765625, 3509.912109, 3940.5761718, 4435.8398437, 4974.1699218, 5577.09960938,
                        6266.1621093, 7041.3574218, 7902.6855468, 8871.6796875])
    assert other.num_bins == 29 and other.num_frames == 281


def test_logarithmic_spectrogram_class(sig, fft_freqs):
    """TestLogarithmicSpectrogramClass and its processor, :270-344."""
    from madmom_b200.audio.stft import ShortTimeFourierTransform
    from madmom_b200.audio.spectrogram import (Spectrogram, LogarithmicSpectrogram, SpectrogramDifference,
                                               FilteredSpectrogram, LogarithmicSpectrogramProcessor)
    result = LogarithmicSpectrogram(sig)
    assert isinstance(result, LogarithmicSpectrogram) and isinstance(result, Spectrogram)
    assert isinstance(result.stft, ShortTimeFourierTransform) and isinstance(result.bin_frequencies, np.ndarray)
    assert isinstance(result.mul, float) and isinstance(result.add, float)
    assert isinstance(result.num_frames, int) and isinstance(result.num_bins, int)
    assert np.allclose(result[0, :8], LOG_HEAD) and np.allclose(result[0, -8:], LOG_TAIL)
    assert result.shape == (281, 1024) and np.allclose(result.bin_frequencies, fft_freqs)
    assert result.mul == 1 and result.add == 1 and result.num_frames == 281 and result.num_bins == 1024
    other = LogarithmicSpectrogram(sig, mul=2, add=2)
    assert other.mul == 2 and other.add == 2
    spec = Spectrogram(sig)                                                            # test_from_spec
    assert np.allclose(LogarithmicSpectrogram(spec)[0, :8], LOG_HEAD)
    assert np.allclose(spec[0, :8], SPEC_HEAD)
    assert isinstance(result.diff(), SpectrogramDifference) and isinstance(result.filter(), FilteredSpectrogram)
    proc = LogarithmicSpectrogramProcessor()
    assert isinstance(proc.mul, float) and isinstance(proc.add, float) and proc.mul == 1 and proc.add == 1
    out = proc.process(sig)
    assert isinstance(out, LogarithmicSpectrogram) and out.shape == (281, 1024)


def test_logarithmic_filtered_spectrogram_class(sig, log_centres):
    """TestLogarithmicFilteredSpectrogramClass and its processor, :347-439."""
    from madmom_b200.audio.stft import ShortTimeFourierTransform
    from madmom_b200.audio.filters import Filterbank, LogarithmicFilterbank
    from madmom_b200.audio.spectrogram import (Spectrogram, LogarithmicFilteredSpectrogram,
                                               LogarithmicFilteredSpectrogramProcessor)
    result = LogarithmicFilteredSpectrogram(sig)
    assert isinstance(result, LogarithmicFilteredSpectrogram) and isinstance(result, Spectrogram)
    assert isinstance(result.stft, ShortTimeFourierTransform)
    assert isinstance(result.filterbank, Filterbank) and isinstance(result.filterbank, LogarithmicFilterbank)
    assert isinstance(result.bin_frequencies, np.ndarray)
    assert isinstance(result.mul, float) and isinstance(result.add, float)
    assert isinstance(result.num_frames, int) and isinstance(result.num_bins, int)
    assert np.allclose(result[0, :8], LOGFILT_HEAD) and np.allclose(result[0, -8:], LOGFILT_TAIL)
    assert result.shape == (281, 81) and result.mul == 1 and result.add == 1
    assert np.allclose(result.bin_frequencies, log_centres)
    assert result.num_frames == 281 and result.num_bins == 81
    other = LogarithmicFilteredSpectrogram(sig, mul=2, add=2)
    assert other.mul == 2 and other.add == 2
    spec = Spectrogram(sig)                                                            # test_from_spec
    from_spec = LogarithmicFilteredSpectrogram(spec)
    assert from_spec.shape == (281, 81) and np.allclose(from_spec[0, :8], LOGFILT_HEAD)
    assert spec.shape == (281, 1024) and np.allclose(spec[0, :8], SPEC_HEAD)
    proc = LogarithmicFilteredSpectrogramProcessor()
    assert proc.filterbank == LogarithmicFilterbank
    assert isinstance(proc.num_bands, int) and isinstance(proc.fmin, float) and isinstance(proc.fmax, float)
    assert isinstance(proc.fref, float) and isinstance(proc.norm_filters, bool)
    assert isinstance(proc.unique_filters, bool) and isinstance(proc.mul, float) and isinstance(proc.add, float)
    assert proc.num_bands == 12 and proc.fmin == 30 and proc.fmax == 17000 and proc.fref == 440
    assert proc.norm_filters is True and proc.unique_filters is True and proc.mul == 1 and proc.add == 1
    out = proc.process(sig)
    assert isinstance(out, LogarithmicFilteredSpectrogram) and out.shape == (281, 81)


def test_spectrogram_difference_class(sig, fft_freqs):
    """TestSpectrogramDifferenceClass, :442-480."""
    from madmom_b200.audio.spectrogram import Spectrogram, SpectrogramDifference
    result = SpectrogramDifference(sig)
    assert isinstance(result, SpectrogramDifference) and isinstance(result, Spectrogram)
    assert isinstance(result.spectrogram, Spectrogram) and isinstance(result.bin_frequencies, np.ndarray)
    assert isinstance(result.diff_ratio, float) and isinstance(result.diff_frames, int)
    assert result.diff_max_bins is None and isinstance(result.positive_diffs, bool)
    assert isinstance(result.num_frames, int) and isinstance(result.num_bins, int)
    assert np.allclose(result[1, :8], [1.13179708, -1.1511457, 2.7810955, 2.39441729, -4.87367058, -0.90269375,
                                       3.48209763, 11.14723015])
    assert np.allclose(result[1, -8:], [-0.01463442, -0.01408007, -0.01462659, -0.01431422, -0.01404046,
                                        -0.01457103, -0.01443923, -0.01443416])
    assert result.shape == (281, 1024) and np.allclose(result.bin_frequencies, fft_freqs)
    assert result.diff_ratio == 0.5 and result.diff_frames == 1 and result.diff_max_bins is None
    assert result.positive_diffs is False and result.num_bins == 1024 and result.num_frames == 281
    assert result.positive_diff().min() == 0


def test_spectrogram_difference_processor_class(sig):
    """TestSpectrogramDifferenceProcessorClass, :483-573.  Continuing a stream frame by frame (`reset=False`) is the
    one thing of that class the GPU path does not offer: it must say so, not return something else."""
    from madmom_b200.audio.spectrogram import Spectrogram, SpectrogramDifferenceProcessor
    sample_spec = Spectrogram(sig)
    spec = np.asarray(sample_spec)
    proc = SpectrogramDifferenceProcessor()
    assert isinstance(proc.diff_ratio, float) and proc.diff_frames is None and proc.diff_max_bins is None
    assert isinstance(proc.positive_diffs, bool) and proc.stack_diffs is None
    assert proc.diff_ratio == 0.5 and proc.positive_diffs is False
    result = proc.process(sample_spec)
    assert result.shape == (281, 1024)
    res = np.asarray(result)
    assert np.sum(res[:1]) == 0 and np.max(res[:2]) > 0 and np.min(res) < 0 and np.allclose(res[0], 0)
    assert np.allclose(res[1:], np.diff(spec, axis=0))
    assert np.allclose(np.asarray(proc.process(sample_spec)), res)                      # a second call: the same
    with pytest.raises(NotImplementedError):
        proc.process(sample_spec, reset=False)
    proc = SpectrogramDifferenceProcessor(diff_frames=2)                               # test_diff_frames
    assert proc.diff_frames == 2
    result = proc.process(sample_spec)
    res = np.asarray(result)
    assert result.shape == (281, 1024) and np.sum(res[:2]) == 0 and np.min(res) < 0 and result.diff_frames == 2
    assert np.allclose(np.asarray(proc.process(sample_spec)), res)
    proc.reset()
    assert proc.diff_frames == 2
    proc = SpectrogramDifferenceProcessor(diff_frames=2)                               # test_positive_diffs
    proc.positive_diffs = True
    res = np.asarray(proc.process(sample_spec))
    assert res.shape == (281, 1024) and np.sum(res[:2]) == 0 and np.min(res) >= 0
    for stack, shape in ((np.hstack, (281, 1024 * 2)), (np.vstack, (281 * 2, 1024)), (np.dstack, (281, 1024, 2))):
        proc.stack_diffs = stack
        res = np.asarray(proc.process(sample_spec))
        assert res.shape == shape and np.min(res) >= 0


def test_superflux_processor_class(sig):
    """TestSuperFluxProcessorClass, :576-599."""
    from madmom_b200.audio.spectrogram import SuperFluxProcessor, SpectrogramDifference
    result = SuperFluxProcessor().process(sig)
    assert np.allclose(result[1, :8], [0.11168772, 0.12317812, 0, 0, 0.03797626, 0.18899226, 0, 0.0903399])
    assert np.allclose(result[1, -8:], [0, 0, 0.01419619, 0, 0.02666602, 0.04325962, 0.10899737, 0.06546581])
    assert isinstance(result, SpectrogramDifference)
    assert result.num_bins == 140 and result.num_frames == 281 and result.shape == (281, 140)
    assert result.diff_ratio == 0.5 and result.diff_max_bins == 3 and result.positive_diffs is True


def test_stages_in_any_order(sig):
    """The reference's classes take any Spectrogram and apply their stage to its data, whatever it went through
    before (`LogarithmicSpectrogram(...).filter()`, tests/test_audio_spectrogram.py:321-324).  Orders the fused
    kernel does not have are computed from the source's device matrix: compared with numpy doing what the
    reference's classes do (np.dot, log10(mul x + add), the lagged difference with zero lead-in rows)."""
    from madmom_b200.audio.spectrogram import (Spectrogram, FilteredSpectrogram, LogarithmicSpectrogram,
                                               LogarithmicFilteredSpectrogram, SpectrogramDifference)
    from madmom_b200.features.onsets import complex_flux
    log_spec = LogarithmicSpectrogram(sig)                                   # (281, 1024), fused
    filt_of_log = log_spec.filter()                                          # filter of a logarithm: generic
    assert isinstance(filt_of_log, FilteredSpectrogram) and filt_of_log.shape == (281, 81)
    fb = np.asarray(filt_of_log.filterbank, dtype=np.float32)
    want = np.dot(np.asarray(log_spec), fb)
    assert np.allclose(np.asarray(filt_of_log), want, rtol=1e-5, atol=1e-6)
    assert np.allclose(filt_of_log.bin_frequencies, FilteredSpectrogram(sig).bin_frequencies)
    # everything built on it stays on the generic path
    log_again = filt_of_log.log(mul=2, add=1)
    assert isinstance(log_again, LogarithmicSpectrogram) and log_again.shape == (281, 81)
    assert np.allclose(np.asarray(log_again), np.log10(2 * want + 1), rtol=1e-5, atol=1e-6)
    d = log_again.diff(diff_frames=2, positive_diffs=True)
    ref = np.zeros_like(want)
    ref[2:] = np.maximum(np.log10(2 * want + 1)[2:] - np.log10(2 * want + 1)[:-2], 0)
    assert isinstance(d, SpectrogramDifference) and np.allclose(np.asarray(d), ref, rtol=1e-5, atol=2e-6)
    assert np.allclose(np.asarray(d.flux()), ref.sum(axis=1), rtol=1e-4, atol=1e-5)
    # the logarithm of a logarithm, the difference of a difference, a filtered filtered spectrogram
    lf = LogarithmicFilteredSpectrogram(sig)
    assert np.allclose(np.asarray(LogarithmicSpectrogram(lf)), np.log10(np.asarray(lf) + 1), rtol=1e-5, atol=1e-6)
    dd = SpectrogramDifference(SpectrogramDifference(lf, diff_frames=1), diff_frames=1)
    first = np.zeros((281, 81), np.float32)
    first[1:] = np.diff(np.asarray(lf), axis=0)
    second = np.zeros_like(first)
    second[1:] = np.diff(first, axis=0)
    assert np.allclose(np.asarray(dd), second, rtol=1e-5, atol=2e-6)
    ff = FilteredSpectrogram(FilteredSpectrogram(sig), num_bands=3)         # a bank over the 81 band centres
    assert ff.shape[0] == 281 and ff.shape[1] == ff.filterbank.shape[1] and ff.filterbank.shape[0] == 81
    assert np.allclose(np.asarray(ff), np.dot(np.asarray(FilteredSpectrogram(sig)),
                                              np.asarray(ff.filterbank, dtype=np.float32)), rtol=1e-5, atol=1e-5)
    assert np.allclose(np.asarray(Spectrogram(filt_of_log)), want, rtol=1e-5, atol=1e-6)     # re-typing keeps it
    with pytest.raises(NotImplementedError):
        complex_flux(filt_of_log)


def test_peak_picking_and_processor_classes(ref_acts):
    """TestPeakPickingFunction.test_online and TestOnsetPeakPickingProcessorClass (tests/test_features_onsets.py:
    247-297) on the reference's own activation fixtures; integers and onset times equal the real reference's
    (tests/golden/make_golden_acts.py).  Online (frame-by-frame) mode is out of scope and says so."""
    from madmom_b200.features.onsets import peak_picking, OnsetPeakPickingProcessor
    act, fps = ref_acts["rnn_act"], float(ref_acts["rnn_fps"])
    onsets = peak_picking(act, threshold=0.23, post_max=0)
    assert list(onsets) == [1, 3, 10, 12, 29, 46, 62, 63, 77, 79, 81, 99, 100, 113, 115, 148, 149, 164, 181, 183,
                            216, 234, 250, 268] and len(onsets) == 24
    assert np.array_equal(onsets, ref_acts["rnn_peaks_online_params"])
    sf, sf_fps = ref_acts["sf_act"], float(ref_acts["sf_fps"])
    proc = OnsetPeakPickingProcessor(threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                                     combine=0.03, delay=0, fps=sf_fps)
    got = proc(sf)
    assert np.allclose(got, [0.01, 0.085, 0.275, 0.445, 0.61, 0.795, 0.98, 1.115, 1.365, 1.475, 1.62, 1.795, 2.14,
                             2.33, 2.485, 2.665])
    assert np.array_equal(got, ref_acts["sf_onsets"])
    proc.delay = 1                                                                       # test_delay
    assert np.allclose(proc(sf) - 1, ref_acts["sf_onsets"]) and np.array_equal(proc(sf), ref_acts["sf_onsets_delay1"])
    with pytest.raises(NotImplementedError):
        OnsetPeakPickingProcessor(threshold=0.23, online=True, fps=fps)
    # what the online processor returns for the whole activation is what the offline one returns with the
    # parameters online mode forces (smooth = post_avg = post_max = 0)
    off = OnsetPeakPickingProcessor(threshold=0.23, smooth=0, post_avg=0, post_max=0, fps=fps)(act)
    assert np.array_equal(off, ref_acts["rnn_onsets_online"])
    for combine in (0.0, 0.03, 0.1):
        p = OnsetPeakPickingProcessor(threshold=0.02, smooth=0.02, pre_avg=0.1, post_avg=0.05, pre_max=0.02,
                                      post_max=0.03, combine=combine, delay=0.005, fps=fps)
        assert np.array_equal(p(act), ref_acts["rnn_onsets_combine_%g" % combine]), combine


def test_spectral_onset_processor_class(sig):
    """TestSpectralOnsetProcessorClass, tests/test_features_onsets.py:155-207: which processors the chain is made of
    for every option, the unknown method, the literal values of the default chain."""
    from madmom_b200.audio.signal import SignalProcessor, FramedSignalProcessor
    from madmom_b200.audio.stft import ShortTimeFourierTransformProcessor
    from madmom_b200.audio.filters import LogarithmicFilterbank
    from madmom_b200.audio.spectrogram import (SpectrogramProcessor, FilteredSpectrogramProcessor,
                                               LogarithmicSpectrogramProcessor)
    from madmom_b200.features.onsets import SpectralOnsetProcessor, spectral_flux, phase_deviation
    proc = SpectralOnsetProcessor()
    assert isinstance(proc.processors[0], SignalProcessor) and isinstance(proc.processors[1], FramedSignalProcessor)
    assert isinstance(proc.processors[2], ShortTimeFourierTransformProcessor)
    assert isinstance(proc.processors[3], SpectrogramProcessor) and proc.processors[4] == spectral_flux
    proc = SpectralOnsetProcessor(filterbank=LogarithmicFilterbank)
    assert isinstance(proc.processors[4], FilteredSpectrogramProcessor) and proc.processors[5] == spectral_flux
    proc = SpectralOnsetProcessor(log=np.log10)
    assert isinstance(proc.processors[4], LogarithmicSpectrogramProcessor) and proc.processors[5] == spectral_flux
    proc = SpectralOnsetProcessor(filterbank=LogarithmicFilterbank, log=np.log10)
    assert isinstance(proc.processors[4], FilteredSpectrogramProcessor)
    assert isinstance(proc.processors[5], LogarithmicSpectrogramProcessor) and proc.processors[6] == spectral_flux
    proc = SpectralOnsetProcessor(onset_method="phase_deviation")
    assert isinstance(proc.processors[2], ShortTimeFourierTransformProcessor)
    assert proc.processors[2].circular_shift and proc.processors[4] == phase_deviation
    with pytest.raises(ValueError):
        SpectralOnsetProcessor(onset_method="nonexistent")
    odf = SpectralOnsetProcessor()(sig)
    assert np.allclose(odf[:6], [0., 100.90120697, 74.44419861, 40.277565, 57.95736313, 46.15561295])


def test_onset_functions_on_the_reference_inputs(sig):
    """The function tests of tests/test_features_onsets.py:35-152 on the same two inputs (:25-27): a circularly
    shifted Spectrogram and LogarithmicFilteredSpectrogram(num_bands=24, mul=1, add=1) of it; literals within
    the stated bound 1e-5 + 1e-4 |ref|, ValueError where the reference raises it."""
    from madmom_b200.audio.spectrogram import Spectrogram, LogarithmicFilteredSpectrogram
    from madmom_b200.features import onsets as ON
    sample_spec = Spectrogram(sig, circular_shift=True)
    lfs = LogarithmicFilteredSpectrogram(sample_spec, num_bands=24, mul=1, add=1)
    for fn, arg, want in (
            (ON.high_frequency_content, lfs, [8.97001563, 9.36399107, 8.64144536, 8.34977449, 8.21097918, 8.40412515]),
            (ON.spectral_diff, lfs, [0, 0.55715936, 0.64004618, 0.0810971, 0.295396, 0.16324584]),
            (ON.spectral_flux, lfs, [0, 3.91207361, 2.91675663, 1.38361311, 2.59582925, 2.16986609]),
            (ON.superflux, lfs, [0, 2.08680153, 0.6411702, 0.38634294, 0.40202433, 0.63349575]),
            (ON.complex_flux, lfs, [0, 0.476213485, 0.0877621323, 0.0593151376, 0.0654867291, 0.0954693183]),
            (ON.modified_kullback_leibler, lfs, [0, 0.71910584, 0.6664055, 0.68092251, 0.69984031, 0.71744561]),
            (ON.phase_deviation, lfs, [0, 0, 0.71957183, 0.91994524, 0.9418999, 0.86083585]),
            (ON.weighted_phase_deviation, sample_spec, [0, 0, 0.19568817, 0.20483065, 0.17890805, 0.16970603]),
            (ON.normalized_weighted_phase_deviation, sample_spec,
             [0, 0, 0.46018526, 0.50193471, 0.42031503, 0.40806249]),
            (ON.complex_domain, sample_spec,
             [399.29980469, 585.9564209, 262.08010864, 225.84718323, 196.88954163, 200.32469177]),
            (ON.rectified_complex_domain, sample_spec, [0, 394.165222, 119.79425, 96.70564, 122.52311, 92.61698])):
        got = np.asarray(fn(arg))[:6]
        assert np.allclose(got, want, rtol=1e-4, atol=1e-5), (fn.__name__, got, want)
    for fn in (ON.weighted_phase_deviation, ON.normalized_weighted_phase_deviation, ON.complex_domain,
               ON.rectified_complex_domain):
        with pytest.raises(ValueError):
            fn(lfs)


def test_stft_processor_phase_and_lgd_classes(sig):
    """ShortTimeFourierTransformProcessorClass, PhaseClass, LocalGroupDelayClass, tests/test_audio_stft.py:218-320."""
    from madmom_b200.audio.stft import (ShortTimeFourierTransform, ShortTimeFourierTransformProcessor, Phase,
                                        LocalGroupDelay, STFT, fft_frequencies)
    proc = ShortTimeFourierTransformProcessor()
    assert proc.window == np.hanning and proc.fft_size is None and proc.circular_shift is False
    result = proc.process(sig)
    assert isinstance(result, ShortTimeFourierTransform) and result.shape == (281, 1024)
    assert np.allclose(result.bin_frequencies, fft_frequencies(1024, 44100)) and result.fft_size == 2048
    assert np.allclose(result.fft_window, np.hanning(2048) / 32767)
    assert result.num_bins == 1024 and result.num_frames == 281
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("error")                    # from a signal the STFT is circularly shifted: no warning
        phase = Phase(sig)
        lgd = LocalGroupDelay(sig)
    assert phase.stft.circular_shift is True
    assert isinstance(phase, Phase) and isinstance(phase.stft, ShortTimeFourierTransform)
    assert isinstance(phase.bin_frequencies, np.ndarray)
    assert isinstance(phase.num_bins, int) and isinstance(phase.num_frames, int)
    assert phase.shape == (281, 1024) and np.allclose(phase.bin_frequencies, fft_frequencies(1024, 44100))
    assert phase.num_bins == 1024 and phase.num_frames == 281
    assert isinstance(phase.local_group_delay(), LocalGroupDelay) and isinstance(phase.lgd(), LocalGroupDelay)
    with pytest.warns(RuntimeWarning):
        Phase(STFT(sig))
    assert isinstance(lgd, LocalGroupDelay) and isinstance(lgd.phase, Phase)
    assert isinstance(lgd.stft, ShortTimeFourierTransform) and isinstance(lgd.bin_frequencies, np.ndarray)
    assert isinstance(lgd.num_bins, int) and isinstance(lgd.num_frames, int)
    assert lgd.shape == (281, 1024) and np.allclose(lgd.bin_frequencies, fft_frequencies(1024, 44100))
    assert lgd.num_bins == 1024 and lgd.num_frames == 281
