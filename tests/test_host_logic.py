"""CPU tests of the host side of the product: integer framing logic, filterbank construction, the processor
plumbing, the lazy result objects (metadata only, nothing is computed without a GPU) and the C-ABI library's
export table.  No compute call is made here."""
import ctypes
import os
import pickle
import re

import numpy as np
import pytest

from oracle import frontend as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ------------------------------------------------------------------ C ABI
def test_library_exports_every_declared_symbol():
    from madmom_b200 import _lib
    from madmom_b200.build import build_library
    build_library()
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "madmom_b200.h")).read()
    declared = set(re.findall(r"MMB_API\s+[\w\s\*]+?\b(mmb_\w+)\s*\(", header))
    assert declared, "no declarations found in the header"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.mmb_version() >= 100
    # struct layouts of the binding match the header (field order / sizes)
    assert ctypes.sizeof(_lib.Track) == 48 and _lib.TRACK_DTYPE.itemsize == 48
    assert ctypes.sizeof(_lib.Outputs) == 13 * 8


def test_framing_helpers_in_library(ref_framing):
    from madmom_b200 import _lib
    for n, hop, end, want in ref_framing["num_frames_cases"]:
        assert _lib.num_frames(int(n), hop, "extend" if end else "normal") == int(want)
    for i in (0, 1, 3, 560, 3599999):
        for hop in (441., 220.5, 441.3):
            assert _lib.frame_start(i, 2048, hop, 7) == O.frame_start(i, 2048, hop, 7)
    with pytest.raises(ValueError):
        _lib.num_frames(10, 441., "bogus")


# ------------------------------------------------------------------ recipe / framing / filters vs the pinned oracle
def test_recipe_scalars(ref_framing):
    from madmom_b200 import recipe as R
    for N, hop, ratio, want in ref_framing["diff_frames_cases"]:
        assert R.diff_frames(ratio, hop, int(N)) == int(want)
    for n, hop, end, want in ref_framing["num_frames_cases"]:
        assert R.num_frames(int(n), hop, "extend" if end else "normal") == int(want)
    assert [R.resolve_origin(o, 2048) for o in ("offline", "online", "stream", 5)] == [0, 1023, -1024, 5]
    assert np.array_equal(R.fft_window(np.hanning, 2048, np.int16), O.fft_window(np.hanning, 2048, np.int16))
    assert np.array_equal(R.fft_window(None, 16, np.int16), np.ones(16) / 32767.)
    assert np.array_equal(R.fft_window(None, 16, np.float32), np.ones(16))
    assert R.SUPERFLUX.resolved_diff_frames() == 1 and R.SUPERFLUX_200.resolved_diff_frames() == 2
    assert [r.resolved_diff_frames() for r in R.RNN_ONSET] == [1, 2, 3]
    assert [r.resolved_diff_frames() for r in R.RNN_BEAT] == [1, 1, 2]
    assert [r.build_filterbank().shape[1] for r in R.RNN_BEAT] == [39, 45, 49]


@pytest.mark.parametrize("N", [1024, 2048, 4096])
@pytest.mark.parametrize("bpo,norm", [(6, True), (12, False), (12, True), (24, False)])
def test_product_filterbank_bit_exact(ref_filterbanks, N, bpo, norm):
    from madmom_b200.audio.filters import LogarithmicFilterbank
    from madmom_b200.recipe import fft_frequencies
    fb = LogarithmicFilterbank(fft_frequencies(N // 2, 44100), bpo, norm_filters=norm)
    key = "N%d_b%d_%s" % (N, bpo, "norm" if norm else "raw")
    s, l, w = fb.band_table()
    assert np.array_equal(s, ref_filterbanks[key + "_start"])
    assert np.array_equal(l, ref_filterbanks[key + "_len"])
    assert np.array_equal(w, ref_filterbanks[key + "_w"])
    assert np.array_equal(fb.center_frequencies, ref_filterbanks[key + "_center"])
    back = pickle.loads(pickle.dumps(fb))
    assert np.array_equal(back, fb) and np.array_equal(back.bin_frequencies, fb.bin_frequencies)


def test_product_filterbank_misc(ref_filterbanks):
    from madmom_b200.audio import filters as F
    fb = F.LogarithmicFilterbank(O.fft_frequencies(1024, 44100), 12, unique_filters=False)
    assert np.array_equal(np.asarray(fb), ref_filterbanks["N2048_b12_norm_nonunique_dense"])
    # reference literals: tests/test_audio_filters.py:224-233, 396-421, 1051-1074
    assert len(F.log_frequencies(6, 30, 17000)) == 55
    assert np.allclose(F.log_frequencies(1, 100, 1000, 441), [110.25, 220.5, 441., 882.])
    assert np.allclose(F.frequencies2bins([0, 1, 2, 3, 4, 5], [0, 2, 4]), [0, 1, 1, 2, 2, 2])
    assert np.allclose(F.frequencies2bins([0, 1, 2, 3, 4, 5], [0, 2, 4], unique_bins=True), [0, 1, 2])
    fb = F.LogarithmicFilterbank(np.arange(0, 20000, 20), 12)
    assert fb.shape == (1000, 81) and fb.num_bands == 81 and fb.num_bins == 1000
    assert np.allclose(fb.min(), 0) and np.allclose(fb.max(), 1)
    mel = F.MelFilterbank(O.fft_frequencies(1024, 44100), 40)
    assert mel.shape[0] == 1024 and mel.shape[1] <= 40
    with pytest.raises(TypeError):
        F.Filterbank(np.zeros(10), np.arange(10))
    with pytest.raises(ValueError):
        F.Filterbank(np.zeros((10, 2)), np.arange(9))


def test_mel_filterbank_bit_exact(ref_mel):
    """MelFilterbank against tables computed by the real reference (filters.py:1035-1092;
    tests/test_audio_filters.py:845-905): the 40-band default and the CNNOnsetProcessor bank (80 bands, 27.5 Hz ..
    16 kHz, normalised, duplicate low bands kept) for all three frame sizes, weights bit for bit."""
    from madmom_b200.audio.filters import MelFilterbank
    from madmom_b200.recipe import fft_frequencies
    fb = MelFilterbank(fft_frequencies(1024, 44100))
    s, l, w = fb.band_table()
    assert fb.shape == (1024, 40)
    assert np.array_equal(s, ref_mel["mel40_N2048_start"]) and np.array_equal(l, ref_mel["mel40_N2048_len"])
    assert np.array_equal(w, ref_mel["mel40_N2048_w"])
    assert np.array_equal(fb.center_frequencies, ref_mel["mel40_N2048_center"])
    for N in (1024, 2048, 4096):
        fb = MelFilterbank(fft_frequencies(N // 2, 44100), num_bands=80, fmin=27.5, fmax=16000, norm_filters=True,
                           unique_filters=False)
        assert fb.shape == (N // 2, 80) and fb.dtype == np.float32
        assert np.array_equal(np.asarray(fb), ref_mel["mel80_N%d_dense" % N])
        assert np.array_equal(fb.center_frequencies, ref_mel["mel80_N%d_center" % N])


def test_gain_attenuate_normalize(ref_mel):
    """adjust_gain / attenuate / normalize (madmom/audio/signal.py:71-167) against the real
    reference's outputs, through the functions and through Signal(norm=..., gain=...)."""
    from madmom_b200.audio import signal as S
    x16, xf = ref_mel["ingest_i16"], ref_mel["ingest_f32"]
    for tag, x in (("i16", x16), ("f32", xf)):
        for gain in (-6.5, 0.0, 3.0):
            key = "gain_%s_%g" % (tag, gain)
            if key not in ref_mel:
                with pytest.raises(ValueError):
                    S.adjust_gain(x, gain)
                continue
            got = S.adjust_gain(x, gain)
            assert got.dtype == x.dtype and np.array_equal(got, ref_mel[key])
        assert np.array_equal(S.attenuate(x, 10.), ref_mel["atten_%s_10" % tag])
        assert S.attenuate(x, 0) is x
    assert np.array_equal(S.normalize(xf), ref_mel["norm_f32"]) and np.array_equal(S.normalize(x16), ref_mel["norm_i16"])
    with pytest.raises(ValueError):
        S.normalize(x16.astype(np.int8))
    sig = S.Signal(x16, sample_rate=44100, norm=True, gain=-6.5)
    assert isinstance(sig, S.Signal) and sig.dtype == np.int16 and sig.sample_rate == 44100
    assert np.array_equal(np.asarray(sig), S.adjust_gain(S.normalize(x16), -6.5))
    proc = S.SignalProcessor(sample_rate=44100, gain=-6.5)
    assert np.array_equal(np.asarray(proc(xf)), ref_mel["gain_f32_-6.5"])


def test_signal_and_framed_signal(sample_pcm, ref_sample):
    from madmom_b200.audio.signal import Signal, FramedSignal, FramedSignalProcessor
    sig = Signal(sample_pcm, sample_rate=44100)
    assert sig.num_samples == 123481 and sig.num_channels == 1 and sig.sample_rate == 44100
    assert np.isclose(sig.length, 2.8)
    frames = FramedSignal(sig)
    assert frames.shape == (281, 2048) and len(frames) == 281 and frames.ndim == 2
    assert frames.hop_size == 441. and frames.fps == 100. and frames.origin == 0
    for r, want in zip(ref_sample["rows"], ref_sample["frames_rows"]):
        assert np.array_equal(frames[int(r)], want)
    assert np.array_equal(frames[-1], frames[280])
    with pytest.raises(IndexError):
        frames[281]
    with pytest.raises(ValueError):
        frames[::2]
    with pytest.raises(TypeError):
        frames["a"]
    f200 = FramedSignalProcessor(frame_size=2048, fps=200)(sig)
    assert f200.num_frames == 561 and f200.hop_size == 220.5
    sl = f200[3:10]
    assert sl.num_frames == 7 and sl.origin == int(-220.5 * 3)
    assert FramedSignal(sig, origin="online").origin == 1023 and FramedSignal(sig, origin="stream").origin == -1024
    assert FramedSignal(sig, end="extend").num_frames == O.num_frames(len(sample_pcm), 441., "extend")
    assert FramedSignalProcessor(fps=100)(sig, fps=50).hop_size == 882.        # process() kwargs override
    back = pickle.loads(pickle.dumps(frames))
    assert back.num_frames == 281 and np.array_equal(np.asarray(back.signal), sample_pcm)
    stereo = Signal(np.stack((sample_pcm, sample_pcm // 2), axis=1), sample_rate=44100)
    assert stereo.num_channels == 2 and FramedSignal(stereo).shape == (281, 2048, 2)
    mono = Signal(stereo, num_channels=1)
    assert np.array_equal(np.asarray(mono), np.mean(np.asarray(stereo), axis=-1).astype(np.int16))


def test_processor_plumbing(tmp_path):
    from madmom_b200.processors import Processor, SequentialProcessor, ParallelProcessor

    class Add(Processor):
        def __init__(self, k):
            self.k = k

        def process(self, data, scale=1, **kwargs):
            return data + self.k * scale

    seq = SequentialProcessor([Add(1), [Add(2), lambda x: x * 10], None])
    assert isinstance(seq[1], SequentialProcessor) and len(seq) == 3
    assert seq(1) == 40 and seq(1, scale=2) == 70                  # kwargs reach Processors, not bare callables
    par = ParallelProcessor([Add(1), Add(2)], num_threads=4)
    assert par(1) == [2, 3]
    assert SequentialProcessor((par, sum))(1) == 5
    seq.append(Add(5))
    assert seq(1) == 45
    path = str(tmp_path / "p.pkl")
    from madmom_b200.audio.signal import FramedSignalProcessor
    FramedSignalProcessor(frame_size=1024, fps=50).dump(path)
    back = Processor.load(path)
    assert isinstance(back, FramedSignalProcessor) and back.frame_size == 1024 and back.fps == 50


def test_lazy_results_metadata_without_gpu(sample_pcm):
    """Constructing the whole chain allocates nothing and needs no GPU; the attribute graph the reference's
    consumers walk is there (spectrogram.py:1114-1118, onsets.py:236)."""
    from madmom_b200.audio.signal import Signal, FramedSignalProcessor
    from madmom_b200.audio.spectrogram import (SuperFluxProcessor, SpectrogramDifferenceProcessor,
                                               LogarithmicFilteredSpectrogramProcessor, MultiFrameProcessor)
    from madmom_b200.features.onsets import SpectralOnsetProcessor, superflux
    sig = Signal(sample_pcm, sample_rate=44100)
    res = SuperFluxProcessor()(FramedSignalProcessor()(sig))
    assert res.shape == (281, 140) and res.diff_frames == 1 and res.positive_diffs and res.diff_max_bins == 3
    spec = res.spectrogram
    assert spec.shape == (281, 140) and spec.mul == 1 and spec.add == 1
    assert spec.filterbank.shape == (1024, 140) and len(spec.bin_frequencies) == 140
    assert spec.stft.frames.hop_size == 441. and spec.stft.frames.frame_size == 2048
    assert spec.stft.frames.signal.sample_rate == 44100 and spec.stft.window.shape == (2048,)
    assert spec.stft.shape == (281, 1024) and len(spec.stft.bin_frequencies) == 1024
    assert np.array_equal(spec.stft.fft_window, O.fft_window(np.hanning, 2048, np.int16))
    assert res._tensor is None and "lazy" in repr(res)
    lf = LogarithmicFilteredSpectrogramProcessor(num_bands=6, mul=5)(FramedSignalProcessor(frame_size=4096)(sig))
    st = SpectrogramDifferenceProcessor(diff_ratio=0.25, positive_diffs=True, stack_diffs=np.hstack)(lf)
    assert st.shape == (281, 98) and st.diff_frames == 3
    proc = SpectralOnsetProcessor(onset_method="superflux", fps=200, num_bands=24, log=np.log10)
    assert proc.processors[-1] is superflux
    with pytest.raises(ValueError):
        SpectralOnsetProcessor(onset_method="nope")
    for p in (SuperFluxProcessor(), MultiFrameProcessor(), proc, SpectrogramDifferenceProcessor()):
        pickle.loads(pickle.dumps(p))


def test_no_cpu_fallback(sample_pcm, monkeypatch):
    """Without a CUDA device every compute access fails loudly (there is no CPU path to fall back to)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    from madmom_b200 import _lib
    from madmom_b200.audio.signal import Signal, FramedSignal
    from madmom_b200.audio.spectrogram import Spectrogram
    spec = Spectrogram(FramedSignal(Signal(sample_pcm, sample_rate=44100)))
    with pytest.raises(_lib.LibraryError):
        np.asarray(spec)
    monkeypatch.setattr(_lib, "LIBPATH", "/nonexistent/libmadmom_b200.so")
    monkeypatch.setattr(_lib, "_lib", None)
    with pytest.raises(_lib.LibraryError):
        _lib.load()


def test_product_does_not_import_the_oracle():
    import subprocess
    import sys
    code = ("import sys; import madmom_b200, madmom_b200.batch, madmom_b200.audio.spectrogram, "
            "madmom_b200.features.onsets, madmom_b200.parallel; "
            "bad = [m for m in sys.modules if m == 'oracle' or m.startswith('oracle.') or m.startswith('scipy.fft')]; "
            "sys.exit(1 if bad else 0)")
    assert subprocess.call([sys.executable, "-c", code], cwd=ROOT) == 0


def test_window_cache_and_plan_key():
    """Per-file overheads of the drop-in path: numpy's window functions are built once per frame size (read-only),
    other callables every time; the plan-cache key of a filterbank is a digest kept on the frozen array;
    diff_frames keeps the reference's values (tests/test_audio_spectrogram.py:459-480 via ref_framing)."""
    from madmom_b200 import recipe as R
    from madmom_b200._engine import _filterbank_key
    from madmom_b200.audio.filters import LogarithmicFilterbank
    w = R.window_of(np.hanning, 2048)
    assert w is R.window_of(np.hanning, 2048) and not w.flags.writeable
    assert np.array_equal(w, np.hanning(2048))
    calls = []

    def custom(n):
        calls.append(n)
        return np.ones(n)
    assert R.window_of(custom, 8).shape == (8,) and R.window_of(custom, 8) is not R.window_of(custom, 8)
    assert calls == [8, 8, 8]
    given = np.arange(4.)
    assert R.window_of(given, 4) is given and R.window_of(None, 4) is None
    for ratio, hop, N, want in ((0.5, 441., 2048, 1), (0.5, 220.5, 2048, 2), (0.25, 441., 4096, 3),
                                (0.5, 220.5, 4096, 5), (0.25, 441., 1024, 1)):
        assert R.diff_frames(ratio, hop, N) == want
    fb = LogarithmicFilterbank(R.fft_frequencies(1024, 44100), num_bands=24, norm_filters=False)
    k1 = _filterbank_key(fb)
    assert k1 == _filterbank_key(fb) and not fb.flags.writeable and k1[0] == (1024, 140)
    other = LogarithmicFilterbank(R.fft_frequencies(1024, 44100), num_bands=12, norm_filters=False)
    assert _filterbank_key(other) != k1 and _filterbank_key(None) is None
    with pytest.raises(ValueError):
        fb[0, 0] = 1.0                      # frozen: a changed filterbank cannot run with the old plan
    same = LogarithmicFilterbank(R.fft_frequencies(1024, 44100), num_bands=24, norm_filters=False)
    assert _filterbank_key(same) == k1      # equal weights, another object: same plan


def test_framed_signal_like_the_reference_tests(sample_pcm):
    """The reference's FramedSignal assertions, tests/test_audio_signal.py:1144-1307 (small arrays, `end`, slicing,
    the origin literals, fps), on the drop-in class."""
    from madmom_b200.audio.signal import Signal, FramedSignal
    result = FramedSignal(np.arange(10), 4, 2)
    assert np.allclose(result[0], [0, 0, 0, 1])
    assert result.frame_size == 4 and result.hop_size == 2. and result.origin == 0 and result.num_frames == 5
    assert len(result) == 5 and result.frame_rate is None and result.fps is None
    assert result.overlap_factor == 0.5 and result.shape == (5, 4) and result.ndim == 2
    assert FramedSignal(np.arange(10), 4, 2, end="extend").num_frames == 6
    result = FramedSignal(np.arange(10), 4, 2, sample_rate=4)
    for i, want in enumerate(([0, 0, 0, 1], [0, 1, 2, 3], [2, 3, 4, 5], [4, 5, 6, 7], [6, 7, 8, 9])):
        assert np.allclose(result[i], want)
    assert np.allclose(result[-1], [6, 7, 8, 9])
    with pytest.raises(IndexError):
        result[5]
    assert result.frame_rate == 2 and result.fps == 2 and len(result) == 5
    sl = result[1:]                                                         # test_values_slicing
    assert np.allclose(sl[0], [0, 1, 2, 3]) and np.allclose(sl[-2], [4, 5, 6, 7]) and np.allclose(sl[3], [6, 7, 8, 9])
    with pytest.raises(IndexError):
        sl[4]
    assert sl.frame_size == 4 and sl.hop_size == 2. and sl.origin == -2 and sl.num_frames == 4
    assert len(sl) == 4 and sl.shape == (4, 4) and sl.fps == 2 and sl.overlap_factor == 0.5
    sl = result[2:4]
    assert sl.shape == (2, 4) and np.allclose(sl[0], [2, 3, 4, 5]) and np.allclose(sl[1], [4, 5, 6, 7])
    with pytest.raises(IndexError):
        sl[2]
    with pytest.raises(ValueError):
        result[2:4:2]
    with pytest.raises(TypeError):
        result["foo":"bar"]
    with pytest.raises(TypeError):
        result["bar"]
    sig = Signal(sample_pcm, sample_rate=44100)                             # test_values_file and the origins
    frames = FramedSignal(sig)
    assert np.allclose(frames[0][:5], [0, 0, 0, 0, 0]) and np.allclose(frames[3], sample_pcm[299: 299 + 2048])
    assert frames.overlap_factor == 0.78466796875 and frames.frame_rate == 100.
    online = FramedSignal(sig, origin="online")
    assert online.origin == 1023 and online.num_frames == 281
    assert np.allclose(online[6], sample_pcm[599: 599 + 2048])
    for literal, want in (("left", 1023), (10, 10), ("offline", 0), ("center", 0), ("right", -1024),
                          ("future", -1024), (-10, -10)):
        assert FramedSignal(sig, origin=literal).origin == want, literal
    shifted = FramedSignal(sig, origin=-10)
    assert shifted.num_frames == 281 and np.allclose(shifted[3], sample_pcm[309: 309 + 2048])
    assert FramedSignal(sig, fps=200).hop_size == 220.5 and FramedSignal(sig, fps=50).hop_size == 882.


def test_signal_frame_and_signal_processor_like_the_reference_tests(sample_pcm, tmp_path):
    """tests/test_audio_signal.py:796-995 on the drop-in `signal_frame` and `SignalProcessor`: frames of short,
    stereo and shifted inputs, a fractional frame size, the processor's defaults, the literal samples after
    normalisation and after -10 dB (int16 truncation), dump / load."""
    from madmom_b200.audio.signal import Signal, SignalProcessor, signal_frame
    from madmom_b200.processors import Processor
    assert signal_frame(np.arange(10), 0, 4, 2).dtype == int
    assert signal_frame(np.arange(10, dtype=float), 0, 4, 2).dtype == float
    sig = Signal(sample_pcm, sample_rate=44100)
    for args in ((0, 4, 2), (2000, 400, 200), (-10, 400, 200)):
        assert np.asarray(signal_frame(sig, *args)).dtype == np.int16
    for idx, want in ((0, [0, 0, 0, 0, 0, 0, 1, 2, 3, 0]), (1, [0, 1, 2, 3, 0, 0, 0, 0, 0, 0]), (2, [0] * 10),
                      (-2, [0] * 10)):
        assert np.allclose(signal_frame(np.arange(4), idx, 10, 5), want)                 # test_short_input_length
    for idx, want in ((-1, [0, 0, 0, 0]), (0, [0, 0, 0, 1]), (1, [0, 1, 2, 3]), (2, [2, 3, 4, 5]), (3, [4, 5, 6, 7]),
                      (4, [6, 7, 8, 9]), (5, [8, 9, 0, 0])):
        assert np.allclose(signal_frame(np.arange(10), idx, 4, 2), want)                 # test_values
    stereo = np.tile(np.arange(10)[:, np.newaxis], 2)
    for idx, want in ((0, [0, 0, 0, 1]), (1, [0, 1, 2, 3]), (5, [8, 9, 0, 0]), (6, [0, 0, 0, 0])):
        assert np.allclose(signal_frame(stereo, idx, 4, 2), np.tile(np.array(want)[:, None], 2))
    for idx, want in ((0, [0, 0, 1]), (1, [1, 2, 3]), (2, [3, 4, 5])):
        assert np.allclose(signal_frame(np.arange(10), idx, 3.5, 2), want)               # test_float_hop_size
    for origin, want in ((1, [0, 0, 0, 0]), (-1, [0, 0, 1, 2]), (-2, [0, 1, 2, 3]), (-4, [2, 3, 4, 5])):
        assert np.allclose(signal_frame(np.arange(10), 0, 4, 2, origin), want)           # test_origin
    assert np.allclose(signal_frame(np.arange(10), 0, 5, 2, -2.5), [0, 1, 2, 3, 4])
    proc = SignalProcessor()
    assert isinstance(proc, Processor)
    assert proc.sample_rate is None and proc.num_channels is None and proc.start is None and proc.stop is None
    assert proc.norm is False and isinstance(proc.gain, float) and proc.gain == 0
    result = proc.process(sig)
    assert isinstance(result, Signal) and isinstance(result, np.ndarray) and result.dtype == np.int16
    assert np.allclose(result[:5], [-2494, -2510, -2484, -2678, -2833]) and len(result) == 123481
    assert result.min() == -20603 and result.max() == 17977 and result.mean() == -172.88385257650975
    assert result.sample_rate == 44100 and result.num_samples == 123481 and result.num_channels == 1
    proc.norm = True                                                                      # test_process_norm
    result = proc.process(sig)
    assert result.dtype == np.int16 and np.allclose(result[:5], [-3966, -3991, -3950, -4259, -4505])
    assert result.min() == -32767 and result.max() == 28590 and result.mean() == -274.92599671204476
    proc.norm, proc.gain = False, -10                                                     # test_process_gain
    result = proc.process(sig)
    assert result.dtype == np.int16 and np.allclose(result[:5], [-788, -793, -785, -846, -895])
    assert len(result) == 123481 and result.sample_rate == 44100
    proc = SignalProcessor()
    proc.dump(str(tmp_path / "signal_processor.pkl"))                                     # test_pickle
    back = SignalProcessor.load(str(tmp_path / "signal_processor.pkl"))
    assert type(back) is SignalProcessor
    for name in ("sample_rate", "num_channels", "start", "stop", "norm", "gain"):
        assert getattr(back, name) == getattr(proc, name)


def test_framed_signal_processor_like_the_reference_tests(sample_pcm, tmp_path):
    """tests/test_audio_signal.py:1338-1440 on the drop-in FramedSignalProcessor: defaults, `origin='online'`, fps,
    `end`, the wrong `end`, dump / load."""
    from madmom_b200.audio.signal import Signal, FramedSignal, FramedSignalProcessor
    from madmom_b200.processors import Processor
    sig = Signal(sample_pcm, sample_rate=44100)
    proc = FramedSignalProcessor()
    assert isinstance(proc, Processor) and isinstance(proc.process(sig), FramedSignal)
    assert proc.frame_size == 2048 and proc.hop_size == 441. and proc.fps is None and proc.origin == 0
    assert proc.end == "normal" and proc.num_frames is None
    result = proc.process(sig)
    assert np.allclose(result[0][:1023], np.zeros(1023)) and np.allclose(result[0][1024], -2494)
    assert result.frame_size == 2048 and result.hop_size == 441. and result.origin == 0 and result.num_frames == 281
    assert len(result) == 281 and result.fps == 100. and result.frame_rate == 100.
    assert result.overlap_factor == 0.78466796875 and result.shape == (281, 2048) and result.ndim == 2
    proc.end = "bogus"
    assert proc.end == "bogus"
    proc.end = "normal"
    proc.origin = "online"                                                    # test_process_online
    result = proc.process(sig)
    assert np.allclose(result[0][-1], -2494) and len(result) == 281 and result.num_frames == 281
    proc.origin = 0
    proc.fps = 200.                                                           # test_process_fps
    result = proc.process(sig)
    assert np.allclose(result[0][:1023], np.zeros(1023)) and np.allclose(result[0][1024], -2494)
    assert len(result) == 561 and result.num_frames == 561
    proc.fps = None
    proc.frame_size, proc.hop_size = 10, 6                                    # test_process_end
    assert proc.process(np.arange(18)).num_frames == 3
    proc.end = "extend"
    assert len(proc.process(np.arange(18))) == 4
    with pytest.raises(ValueError):
        FramedSignalProcessor(end="bla").process(sig)
    proc = FramedSignalProcessor()
    proc.dump(str(tmp_path / "framed.pkl"))
    back = FramedSignalProcessor.load(str(tmp_path / "framed.pkl"))
    assert type(back) is FramedSignalProcessor
    for name in ("frame_size", "hop_size", "fps", "end"):
        assert getattr(back, name) == getattr(proc, name)


def test_filter_functions_and_filterbank_classes_like_the_reference_tests(ref_filterbanks):
    """tests/test_audio_filters.py:206-236 (log_frequencies), :376-448 (frequencies2bins, bins2frequencies),
    :845-921 (MelFilterbank), :1023-1074 (LogarithmicFilterbank) on the drop-in module."""
    from madmom_b200.audio import filters as F
    from madmom_b200.audio.stft import fft_frequencies
    for args in ((), (1,), (1, 2), (1, 2, 3, 4, 5)):
        with pytest.raises(TypeError):
            F.log_frequencies(*args)
    result = F.log_frequencies(12, 20, 20000)
    assert isinstance(result, np.ndarray) and result.dtype == float
    assert len(F.log_frequencies(6, 30, 17000)) == 55
    assert np.allclose(F.log_frequencies(1, 100, 1000), [110., 220., 440., 880.])
    assert np.allclose(F.log_frequencies(1, 100, 1000, 441), [110.25, 220.5, 441., 882.])
    for args in ((), (1,), (1, 2, 3, 4)):
        with pytest.raises(TypeError):
            F.frequencies2bins(*args)
    result = F.frequencies2bins([0, 1, 2, 3, 4], [0, 1, 2, 3, 4])
    assert isinstance(result, np.ndarray) and np.issubdtype(result.dtype, np.integer)
    assert np.allclose(result, [0, 1, 2, 3, 4])
    assert np.allclose(F.frequencies2bins([0, 1, 1.1, 1.2, 2, 3, 3.5], [0, 1, 2, 3, 4]), [0, 1, 1, 1, 2, 3, 4])
    assert np.allclose(F.frequencies2bins([0, 1, 2, 3, 11], [0, 1, 2, 3, 4]), [0, 1, 2, 3, 4])
    assert np.allclose(F.frequencies2bins([-1, 0, 1, 2, 3, 4], [0, 1, 2, 3, 4]), [0, 0, 1, 2, 3, 4])
    assert np.allclose(F.frequencies2bins([0, 1, 2, 3, 4, 5], [0, 2, 4]), [0, 1, 1, 2, 2, 2])
    assert np.allclose(F.frequencies2bins([0, 1, 2, 3, 4, 5], [0, 2, 4], unique_bins=True), [0, 1, 2])
    assert np.allclose(F.frequencies2bins([-1, 0, 1, 2, 3, 4], [0, 1, 2, 3, 4], unique_bins=True), [0, 1, 2, 3, 4])
    for args in ((), (1,), (1, 2, 3)):
        with pytest.raises(TypeError):
            F.bins2frequencies(*args)
    result = F.bins2frequencies([0, 1, 2, 3, 4], [0, 1, 2, 3, 4])
    assert isinstance(result, np.ndarray) and result.dtype == float and np.allclose(result, [0, 1, 2, 3, 4])
    assert np.allclose(F.bins2frequencies([0, 1, 2, 3, 4], [0, 1, 1.1, 1.2, 2, 3, 4]), [0, 1, 1.1, 1.2, 2])
    assert np.allclose(F.bins2frequencies([-1, 0, 1, 2, 3, 4], [0, 1, 2, 3, 4]), [4, 0, 1, 2, 3, 4])
    # LogarithmicFilterbank
    filt = F.LogarithmicFilterbank(np.arange(20000))
    assert isinstance(filt, F.LogarithmicFilterbank) and filt.dtype == F.FILTER_DTYPE
    assert filt.bin_frequencies.dtype == float
    with pytest.raises(NotImplementedError):
        F.LogarithmicFilterbank(np.arange(20000), bands_per_octave=False)
    assert isinstance(F.LogarithmicFilterbank.NUM_BANDS_PER_OCTAVE, int)
    assert F.LogarithmicFilterbank.NUM_BANDS_PER_OCTAVE == 12
    filt = F.LogarithmicFilterbank(np.arange(0, 20000, 20), num_bands=12)
    assert np.allclose(filt.min(), 0) and np.allclose(filt.max(), 1) and filt.shape == (1000, 81)
    filt = F.LogarithmicFilterbank(np.arange(0, 20000, 20), num_bands=12, unique_filters=False)
    assert np.allclose(filt.min(), 0) and np.allclose(filt.max(), 1) and filt.shape == (1000, 108)
    filt = F.LogarithmicFilterbank(fft_frequencies(1024, 44100))
    assert filt.num_bands == 81 and filt.num_bands_per_octave == 12 and filt.num_bins == 1024
    assert np.allclose(filt.min(), 0) and np.allclose(filt.fmin, 43.066406) and np.allclose(filt.fmax, 16731.298828)
    assert np.allclose(filt.center_frequencies, ref_filterbanks["N2048_b12_norm_center"])
    corner = np.asarray(filt.corner_frequencies)
    assert corner.shape == (81, 2) and np.all(corner[:, 0] <= filt.center_frequencies)
    assert np.all(corner[:, 1] >= filt.center_frequencies)
    # MelFilterbank
    filt = F.MelFilterbank(np.arange(20000))
    assert isinstance(filt, F.MelFilterbank) and filt.dtype == F.FILTER_DTYPE and filt.bin_frequencies.dtype == float
    assert isinstance(F.MelFilterbank.FMIN, float) and isinstance(F.MelFilterbank.FMAX, float)
    assert isinstance(F.MelFilterbank.NUM_BANDS, int) and isinstance(F.MelFilterbank.NORM_FILTERS, bool)
    assert isinstance(F.MelFilterbank.UNIQUE_FILTERS, bool)
    assert (F.MelFilterbank.FMIN, F.MelFilterbank.FMAX, F.MelFilterbank.NUM_BANDS) == (20., 17000., 40)
    assert F.MelFilterbank.NORM_FILTERS is True and F.MelFilterbank.UNIQUE_FILTERS is True
    filt = F.MelFilterbank(np.arange(1000) * 20, 10)
    assert filt.num_bands == 10 and filt.num_bins == 1000 and filt.fmin == 40 and filt.fmax == 16980
    assert np.allclose(filt.min(), 0) and np.allclose(filt.max(), 0.0714286)
    assert np.allclose(filt.center_frequencies, [260, 580, 1020, 1600, 2380, 3420, 4820, 6700, 9180, 12520])
    assert np.allclose(filt.corner_frequencies, [[40, 560], [280, 1000], [600, 1580], [1040, 2360], [1620, 3400],
                                                 [2400, 4800], [3440, 6680], [4840, 9160], [6720, 12500],
                                                 [9200, 16980]])
    filt = F.MelFilterbank(np.fft.fftfreq(2048, 1. / 44100)[:1024])
    assert np.allclose(filt.min(), 0) and np.allclose(filt.max(), 1. / 3)
    assert np.allclose(filt.center_frequencies[:4], [86.132812, 150.732422, 215.332031, 279.931640])
    assert np.allclose(filt.corner_frequencies[-1], [14470.312500, 16968.164062])


def test_remix_like_the_reference_tests(sample_pcm):
    """TestMixFunction, tests/test_audio_signal.py:277-361, on the drop-in `remix`: shapes and dtypes of the mono and
    multi-channel cases, channel selection, the integer mean, the unsupported conversion."""
    from madmom_b200.audio.signal import Signal, remix
    sig_1d = np.asarray([0, 0, 1, 0, 0, 1, 0, 0, 1], dtype=float)
    sig_2d = np.asarray([[0, 0, 1, 0, 0, 1, 0, 0, 1], [1, 0, 1, 0, 1, 0, 1, 0, 1]], dtype=float).T
    mono_2d = np.asarray([0.5, 0, 1, 0, 0.5, 0.5, 0.5, 0, 1], dtype=float)
    for x, dt in ((sig_1d, float), (sig_1d.astype(int), int)):
        r = remix(x, 1)
        assert len(r) == len(x) and r.shape == x.shape and r.dtype == dt
        for ch in (2, 3):
            r = remix(x, ch)
            assert len(r) == len(x) and r.shape == (len(x), ch) and r.dtype == dt
    sig = Signal(sample_pcm, sample_rate=44100)
    r = remix(sig, 1)
    assert isinstance(r, Signal) and isinstance(r, np.ndarray) and r.dtype == np.int16
    for x, dt in ((sig_2d, float), (sig_2d.astype(int), int)):
        r = remix(x, 1)
        assert len(r) == len(x) and r.shape == sig_1d.shape and r.dtype == dt
        r = remix(x, 2)
        assert len(r) == len(x) and r.shape == x.shape and r.dtype == dt
    stereo = Signal(np.stack((sample_pcm, sample_pcm // 2), axis=1), sample_rate=44100)
    for ch in (1, 2):
        r = remix(stereo, ch)
        assert isinstance(r, Signal) and isinstance(r, np.ndarray) and r.dtype == np.int16
    with pytest.raises(NotImplementedError):
        remix(sig_2d, 3)
    r = remix(sig_2d, 1, channel=0)                                          # test_channel_selection
    assert r.shape == sig_1d.shape and np.array_equal(r, sig_1d)
    assert np.array_equal(remix(sig_2d, 1, channel=1), sig_2d[:, 1])
    assert np.allclose(remix(sig_1d, 1), sig_1d)                             # test_values
    assert np.allclose(remix(sig_2d.astype(int), 1), mono_2d.astype(int))
    assert np.allclose(remix(sig_2d, 1), mono_2d)
    assert np.allclose(remix(2 * sig_2d.astype(int), 1), 2 * mono_2d)


def test_signal_class_like_the_reference_tests(sample_pcm, tmp_path):
    """TestSignalClass, tests/test_audio_signal.py:696-795, on the drop-in Signal: types and values with and without a
    sample rate, 2-D signals, num_channels, the sample file's samples, write / re-read, energy / rms / spl."""
    from madmom_b200.audio.signal import Signal
    sig_1d = np.asarray([0, 0, 1, 0, 0, 1, 0, 0, 1], dtype=float)
    sig_2d = np.asarray([[0, 0, 1, 0, 0, 1, 0, 0, 1], [1, 0, 1, 0, 1, 0, 1, 0, 1]], dtype=float).T
    r = Signal(sig_1d)
    assert isinstance(r, Signal) and isinstance(r, np.ndarray) and r.dtype == float
    assert r.start is None and r.stop is None and isinstance(r.num_samples, int) and r.sample_rate is None
    assert isinstance(r.num_channels, int) and r.length is None and isinstance(r.ndim, int)
    assert np.allclose(r, sig_1d) and len(r) == 9 and r.num_samples == 9 and r.num_channels == 1
    r = Signal(sig_1d, 1)
    assert r.dtype == float and isinstance(r.sample_rate, int) and isinstance(r.length, float)
    assert np.allclose(r, sig_1d) and r.num_samples == 9 and r.sample_rate == 1 and r.length == 9
    r = Signal(sig_2d, 12.3)
    assert np.allclose(r, sig_2d) and len(r) == 9 and r.num_samples == 9 and r.sample_rate == 12.3
    assert r.num_channels == 2 and r.length == 9 / 12.3 and r.ndim == 2
    r = Signal(sig_2d, sample_rate=1, num_channels=1)
    assert r.shape == (9,) and np.allclose(r, [0.5, 0, 1, 0, 0.5, 0.5, 0.5, 0, 1])
    r = Signal(sample_pcm, sample_rate=44100)
    assert np.allclose(r[:5], [-2494, -2510, -2484, -2678, -2833]) and len(r) == 123481
    assert r.num_samples == 123481 and r.sample_rate == 44100 and r.num_channels == 1 and r.ndim == 1
    assert np.allclose(r.length, 2.8)
    path = str(tmp_path / "sample_copy.wav")                                 # test_write_method
    r.write(path)
    back = Signal(path)
    assert back.dtype == np.int16 and back.sample_rate == 44100 and np.allclose(r, back)
    s = Signal(sig_1d)                                                        # test_methods
    assert np.allclose(s.energy(), 3) and np.allclose(s.rms(), 0.57735026919) and np.allclose(s.spl(), -4.7712125472)
    s = Signal(sig_2d)
    assert np.allclose(s.energy(), 8) and np.allclose(s.root_mean_square(), 2. / 3)
    assert np.allclose(s.sound_pressure_level(), -3.52182518111)


def test_energy_rms_spl_like_the_reference_tests():
    """The frame-wise and silence cases of tests/test_audio_signal.py:600-693 (root_mean_square, sound_pressure_level)."""
    from madmom_b200.audio.signal import FramedSignal, remix, root_mean_square, sound_pressure_level
    sig_1d = np.asarray([0, 0, 1, 0, 0, 1, 0, 0, 1], dtype=float)
    sig_2d = np.asarray([[0, 0, 1, 0, 0, 1, 0, 0, 1], [1, 0, 1, 0, 1, 0, 1, 0, 1]], dtype=float).T
    assert isinstance(root_mean_square(sig_1d), float) and np.allclose(root_mean_square(np.zeros(100)), 0)
    assert np.allclose(root_mean_square(FramedSignal(sig_1d, frame_size=4, hop_size=2)), [0, 0.5, 0.70710678, 0.5, 0.5])
    assert np.allclose(root_mean_square(FramedSignal(sig_2d, frame_size=4, hop_size=2)),
                       [0.35355339, 0.61237244, 0.70710678, 0.61237244, 0.61237244])
    lowest = -np.finfo(float).max
    assert np.allclose(sound_pressure_level(np.zeros(100)), lowest)
    sinus = np.cos(np.linspace(0, 2 * np.pi * 100, 2 * 100 + 1))
    assert np.allclose(sound_pressure_level(sinus), 0.)
    sinus_int16 = (sinus * np.iinfo(np.int16).max).astype(np.int16)
    assert np.allclose(sound_pressure_level(sinus_int16), 0., atol=1e-3)
    assert np.allclose(sound_pressure_level(remix(sinus_int16, 2)), 0., atol=1e-3)
    assert np.allclose(sound_pressure_level(FramedSignal(sig_1d, frame_size=4, hop_size=2)),
                       [lowest, -6.0206, -3.0103, -6.0206, -6.0206])
    assert np.allclose(sound_pressure_level(FramedSignal(sig_2d, frame_size=4, hop_size=2)),
                       [-9.03089987, -4.25968732, -3.01029996, -4.25968732, -4.25968732])


def test_gain_normalize_rescale_trim_like_the_reference_tests(sample_pcm):
    """TestAdjustGainFunction, TestAttenuateFunction, TestNormalizeFunction, TestRescaleFunction, TestTrimFunction
    (tests/test_audio_signal.py:95-276, 447-555) on the drop-in functions: shapes, dtypes, values on the small
    signals, the integer truncation, the errors."""
    from madmom_b200.audio.signal import Signal, adjust_gain, attenuate, normalize, rescale, trim
    sig_1d = np.asarray([0, 0, 1, 0, 0, 1, 0, 0, 1], dtype=float)
    sig_2d = np.asarray([[0, 0, 1, 0, 0, 1, 0, 0, 1], [1, 0, 1, 0, 1, 0, 1, 0, 1]], dtype=float).T
    sig = Signal(sample_pcm, sample_rate=44100)
    for fn in (adjust_gain, attenuate):
        for x in (sig_1d, sig_1d.astype(np.int16), sig_2d, sig_2d.astype(int)):
            r = fn(x, 0)
            assert type(r) == type(x) and r.shape == x.shape and r.dtype == x.dtype and np.allclose(r, x)
        r = fn(sig, 0)
        assert isinstance(r, Signal) and r.dtype == np.int16
    assert np.allclose(adjust_gain(sig_1d, -10), 0.31622777 * sig_1d)
    assert np.allclose(adjust_gain(sig_1d, 10), 3.1622777 * sig_1d)
    assert np.allclose(adjust_gain(sig_1d.astype(int), -5), 0 * sig_1d)
    assert np.allclose(adjust_gain(sig_2d, -3), 0.70794578 * sig_2d)
    assert np.allclose(adjust_gain(sig_2d.astype(np.int16), -1), 0 * sig_2d)
    with pytest.raises(ValueError):
        adjust_gain(sig_2d.astype(np.int16), +60)
    assert np.allclose(attenuate(sig_1d, 10), 0.31622777 * sig_1d)
    assert np.allclose(attenuate(sig_1d, -10), 3.1622777 * sig_1d)
    assert np.allclose(attenuate(sig_1d.astype(int), 5), 0 * sig_1d)
    assert np.allclose(attenuate(sig_2d, 3), 0.70794578 * sig_2d)
    assert np.allclose(attenuate(sig_2d.astype(np.int16), 1), 0 * sig_2d)
    with pytest.raises(ValueError):
        attenuate(sig_2d.astype(np.int16), -10)
    for x in (sig_1d, sig_1d.astype(np.int16), sig_2d, sig_2d.astype(np.int32)):       # normalize
        r = normalize(x)
        assert r.shape == x.shape and r.dtype == x.dtype
    r = normalize(sig)
    assert isinstance(r, Signal) and r.dtype == np.int16
    assert np.allclose(normalize(sig_1d), sig_1d) and np.allclose(normalize(sig_1d * 0.5), sig_1d)
    assert np.max(normalize(sig_1d * 0.5)) == 1
    r = normalize(10 * sig_1d.astype(np.int16))
    assert np.allclose(r, sig_1d * 32767) and np.max(r) == 32767
    r = normalize(sig_2d * 4 - 2)
    assert np.allclose(r, sig_2d * 2 - 1) and r.max() == 1 and r.min() == -1
    r = normalize(3 * sig_2d.astype(np.int32))
    assert np.allclose(r, sig_2d * 2147483647) and np.max(r) == 2147483647
    with pytest.raises(ValueError):
        normalize(sig_2d.astype(np.int64))
    r = rescale(sig_1d, float)                                                          # rescale
    assert r.shape == sig_1d.shape and r.dtype == float and np.allclose(r, sig_1d)
    r = rescale(sig)
    assert isinstance(r, Signal) and r.dtype == np.float32
    assert np.allclose(r[:6], [-0.07611316, -0.07660146, -0.07580798, -0.08172857, -0.08645894, -0.08212531])
    r = rescale(sig_2d, np.float16)
    assert r.shape == sig_2d.shape and r.dtype == np.float16 and np.allclose(r, sig_2d)
    assert rescale(sig, float).dtype == float
    for bad in ((sig_2d, complex), (sig_2d, int)):
        with pytest.raises(ValueError):
            rescale(*bad)
    with pytest.raises(ValueError):
        rescale(np.ones(10, dtype=complex))
    r = trim(sig_1d)                                                                    # trim
    assert type(r) == type(sig_1d) and r.dtype == sig_1d.dtype and np.allclose(r, [1, 0, 0, 1, 0, 0, 1])
    r = trim(sig)
    assert isinstance(r, Signal) and r.ndim == 1 and r.dtype == np.int16
    assert trim(sig_2d).shape == sig_2d.shape                   # leading zeros in one channel only stay
    both = np.tile(np.arange(5), 2).reshape(2, 5).T
    r = trim(both)
    assert r.shape == (4, 2) and np.allclose(r[:, 0], np.arange(1, 5)) and np.allclose(r[:, 1], np.arange(1, 5))
    assert np.allclose(trim(np.arange(5, 0, -1)), [5, 4, 3, 2, 1])


def test_filterbank_base_class_like_the_reference_tests():
    """The parts of TestFilterbankClass (tests/test_audio_filters.py:709-833) that apply to the drop-in Filterbank (it
    has no Filter / from_filters layer: banks are built directly): constructor types and errors, and the derived
    properties on the reference's rectangular and triangular example banks written out as dense matrices."""
    from madmom_b200.audio.filters import Filterbank, FILTER_DTYPE
    filt = Filterbank(np.zeros((100, 10)), np.arange(100))
    assert isinstance(filt, Filterbank) and filt.dtype == FILTER_DTYPE and filt.bin_frequencies.dtype == float
    with pytest.raises(TypeError):
        Filterbank(np.zeros((100, 10)))
    with pytest.raises(TypeError):
        Filterbank(np.zeros(10), np.arange(10))
    with pytest.raises(ValueError):
        Filterbank(np.zeros((100, 10)), np.arange(10))
    rect = np.zeros((100, 4))
    for b, (a, z) in enumerate(((0, 10), (10, 25), (25, 50), (50, 100))):
        rect[a:z, b] = 1
    filt = Filterbank(rect, np.arange(100))
    assert filt.num_bands == 4 and filt.num_bins == 100 and filt.fmin == 0 and filt.fmax == 99
    assert np.allclose(filt.min(), 0) and np.allclose(filt.max(), 1)
    assert np.allclose(filt.bin_frequencies, np.arange(100))
    assert np.allclose(filt.corner_frequencies, [[0, 9], [10, 24], [25, 49], [50, 99]])
    assert np.allclose(filt.center_frequencies, [4, 17, 37, 74])
    tri = np.zeros((100, 4))                                   # TriangularFilter(start, center, stop), norm=False
    for b, (a, c, z) in enumerate(((0, 6, 15), (6, 15, 25), (15, 25, 50), (25, 50, 70))):
        tri[a:c, b] = np.linspace(0, 1, c - a, endpoint=False)
        tri[c:z, b] = np.linspace(1, 0, z - c, endpoint=False)
    filt = Filterbank(tri, np.arange(100))
    assert filt.num_bands == 4 and filt.fmin == 1 and filt.fmax == 69
    assert np.allclose(filt.corner_frequencies, [[1, 14], [7, 24], [16, 49], [26, 69]])
    assert np.allclose(filt.center_frequencies, [6, 15, 25, 50])


def test_smooth_like_the_reference_tests():
    """TestSmoothFunction, tests/test_audio_signal.py:27-93."""
    from madmom_b200.audio.signal import smooth
    sig_1d = np.asarray([0, 0, 1, 0, 0, 1, 0, 0, 1], dtype=float)
    sig_2d = np.asarray([[0, 0, 1, 0, 0, 1, 0, 0, 1], [1, 0, 1, 0, 1, 0, 1, 0, 1]], dtype=float).T
    for x in (sig_1d, sig_2d):
        for k in (None, 3):
            r = smooth(x, k)
            assert type(r) == type(x) and r.shape == x.shape
    for bad in ((np.zeros(9).reshape(3, 3), -1), (np.zeros(9).reshape(3, 3), "bla"), (np.zeros(18).reshape(3, 3, 2), 4)):
        with pytest.raises(ValueError):
            smooth(*bad)
    assert np.allclose(smooth(sig_1d, None), sig_1d) and np.allclose(smooth(sig_1d, 0), sig_1d)
    r3 = [0, 0.08, 1, 0.08, 0.08, 1, 0.08, 0.08, 1]
    r5 = [0.08, 0.54, 1, 0.62, 0.62, 1, 0.62, 0.62, 1]
    r7 = [0.31, 0.77, 1.08, 1.08, 1.08, 1.16, 1.08, 1.08, 1.08]
    assert np.allclose(smooth(sig_1d, 3), r3) and np.allclose(smooth(sig_1d, 5), r5)
    assert np.allclose(smooth(sig_1d, 7), r7)
    assert np.allclose(smooth(sig_1d, np.ones(3)), [0, 1, 1, 1, 1, 1, 1, 1, 1])
    assert np.allclose(smooth(sig_1d, np.ones(4)), [0, 1, 1, 1, 2, 1, 1, 2, 1])
    assert np.allclose(smooth(sig_2d, None), sig_2d)
    assert np.allclose(smooth(sig_2d, 3), np.asarray([r3, [1, 0.16, 1, 0.16, 1, 0.16, 1, 0.16, 1]]).T)
    assert np.allclose(smooth(sig_2d, 5), np.asarray([r5, [1.08, 1.08, 1.16, 1.08, 1.16, 1.08, 1.16, 1.08, 1.08]]).T)
    assert np.allclose(smooth(sig_2d, 7), np.asarray([r7, [1.31, 1.62, 1.62, 1.7, 1.62, 1.7, 1.62, 1.62, 1.31]]).T)


def test_stft_module_functions_like_the_reference_tests():
    """TestBinFrequenciesFunction, TestPhaseFunction, TestLocalGroupDelayFunction (tests/test_audio_stft.py:23-44,
    107-148) on the host utilities of the drop-in module.  The generic `stft()` of the reference (any frame size, on
    the host) is deliberately not offered: a CPU transform in this package would be a CPU fallback."""
    from madmom_b200.audio import stft as ST
    for args in ((), (1,), (1, 2, 3)):
        with pytest.raises(TypeError):
            ST.fft_frequencies(*args)
    r = ST.fft_frequencies(5, 10)
    assert isinstance(r, np.ndarray) and r.dtype == float and np.allclose(r, [0, 1, 2, 3, 4])
    assert not hasattr(ST, "stft")
    rng = np.random.default_rng(0)
    for shape in ((10,), (10, 2)):
        r = ST.phase(rng.random(shape))
        assert r.dtype == float and r.shape == shape
        data = rng.random(shape) + 1j * rng.random(shape)
        r = ST.phase(data)
        assert r.dtype == float and r.shape == shape and np.allclose(np.angle(data), r)
    r = ST.local_group_delay(rng.random((10, 2)))
    assert r.dtype == float and r.shape == (10, 2)
    with pytest.raises(ValueError):
        ST.local_group_delay(np.arange(10))
    with pytest.raises(ValueError):
        ST.local_group_delay(np.arange(20).reshape(5, 2, 2))
    assert np.allclose(np.tile([-2, 0], 10).reshape(10, 2), ST.local_group_delay(np.arange(20).reshape(10, 2) * 2))
    assert np.allclose(np.tile([2.28318531, 0], 10).reshape(10, 2),
                       ST.local_group_delay(np.arange(20).reshape(10, 2) * 4))
    assert ST.lgd is ST.local_group_delay


def test_spec_function_like_the_reference_tests():
    """TestSpecFunction, tests/test_audio_spectrogram.py:44-67."""
    from madmom_b200.audio.spectrogram import spec
    rng = np.random.default_rng(1)
    for shape in ((10,), (10, 2)):
        r = spec(rng.random(shape))
        assert r.dtype == float and r.shape == shape
        data = rng.random(shape) + 1j * rng.random(shape)
        r = spec(data)
        assert r.dtype == float and r.shape == shape and np.allclose(np.abs(data), r)
