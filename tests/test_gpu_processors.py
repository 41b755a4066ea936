"""GPU parity tests of the drop-in processor layer (madmom_b200.audio.*, madmom_b200.features.onsets) against the
reference's golden vectors (tests/golden/, produced by the real madmom) and the CPU oracle.  They read like the
reference's own tests (tests/test_audio_spectrogram.py, tests/test_features_onsets.py).
Bounds: integers bit-exact; float32 features |gpu - ref| <= 1e-5 + 1e-4 |ref|."""
import pickle

import numpy as np
import pytest

from conftest import assert_close

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def sig(sample_pcm):
    from madmom_b200.audio.signal import Signal
    return Signal(sample_pcm, sample_rate=44100)


def test_spectrogram_stages(sig, ref_sample):
    from madmom_b200.audio.signal import FramedSignal
    from madmom_b200.audio.stft import ShortTimeFourierTransform
    from madmom_b200.audio.spectrogram import (Spectrogram, FilteredSpectrogram, LogarithmicSpectrogram,
                                               LogarithmicFilteredSpectrogram)
    rows = ref_sample["rows"]
    frames = FramedSignal(sig, frame_size=2048, hop_size=441.)
    stft = ShortTimeFourierTransform(frames)
    assert stft.shape == (281, 1024) and stft.dtype == np.complex64
    scale = np.abs(ref_sample["stft_rows"]).max()
    assert np.abs(np.asarray(stft)[rows] - ref_sample["stft_rows"]).max() <= 2e-6 * scale
    spec = Spectrogram(stft)
    assert spec.shape == (281, 1024) and spec.num_bins == 1024 and spec.num_frames == 281
    # reference: tests/test_audio_spectrogram.py:82-96
    assert np.allclose(spec[0, :8], [3.15249, 4.00272, 5.66156, 6.30141, 6.02199, 10.84909, 17.83130, 19.44511])
    assert np.abs(np.asarray(spec)[rows] - ref_sample["spec_rows"]).max() <= 2e-6 * scale
    assert np.allclose(np.asarray(spec, dtype=np.float64).sum(axis=0), ref_sample["spec_col_sums"], rtol=1e-4,
                       atol=1e-4)
    filt = FilteredSpectrogram(spec)
    assert filt.shape == (281, 81) and filt.filterbank.shape == (1024, 81)
    assert_close(np.asarray(filt), ref_sample["filt12n"], what="FilteredSpectrogram")
    logfilt = LogarithmicFilteredSpectrogram(spec)
    assert_close(np.asarray(logfilt), ref_sample["logfilt12n"], what="LogarithmicFilteredSpectrogram")
    assert np.allclose(logfilt.bin_frequencies[:3], filt.bin_frequencies[:3])
    # log of the unfiltered spectrogram (tests/test_audio_spectrogram.py:285-293)
    logspec = LogarithmicSpectrogram(spec)
    assert logspec.shape == (281, 1024)
    assert_close(np.asarray(logspec)[rows], ref_sample["logspec_rows"], what="LogarithmicSpectrogram")
    # building from anything a Signal can be made of, kwargs travel down the chain
    lf24 = LogarithmicFilteredSpectrogram(frames, num_bands=24, norm_filters=False)
    assert_close(np.asarray(lf24), ref_sample["logfilt24"], what="logfilt24")


def test_superflux_processor(sig, ref_sample):
    from madmom_b200.audio.signal import FramedSignalProcessor
    from madmom_b200.audio.spectrogram import SuperFluxProcessor, SpectrogramDifference
    proc = SuperFluxProcessor()
    proc = pickle.loads(pickle.dumps(proc))
    res = proc(FramedSignalProcessor()(sig))
    assert isinstance(res, SpectrogramDifference)
    assert res.shape == (281, 140) and res.diff_frames == 1 and res.diff_max_bins == 3
    assert_close(np.asarray(res), ref_sample["superflux_diff"], what="SuperFluxProcessor")
    # tests/test_audio_spectrogram.py:584-591
    assert np.allclose(res[1, :8], [0.11168772, 0.12317812, 0, 0, 0.03797626, 0.18899226, 0, 0.0903399], atol=1e-5)
    assert not np.asarray(res)[0].any()
    again = pickle.loads(pickle.dumps(res))            # results travel through pickle with their host copy
    assert np.array_equal(np.asarray(again), np.asarray(res))
    # keep_dims=False and signed differences
    d2 = SpectrogramDifference(res.spectrogram, diff_frames=2, diff_max_bins=3, positive_diffs=False)
    assert_close(np.asarray(d2), ref_sample["diff_signed_d2"], what="signed d=2")
    d2k = SpectrogramDifference(res.spectrogram, diff_frames=2, diff_max_bins=3, keep_dims=False)
    assert d2k.shape == (279, 140)
    assert np.array_equal(np.asarray(d2k), np.asarray(d2)[2:])


def _multires_chain(mul, ratio):
    from madmom_b200.processors import ParallelProcessor, SequentialProcessor
    from madmom_b200.audio.signal import FramedSignalProcessor
    from madmom_b200.audio.stft import ShortTimeFourierTransformProcessor
    from madmom_b200.audio.spectrogram import (FilteredSpectrogramProcessor, LogarithmicSpectrogramProcessor,
                                               SpectrogramDifferenceProcessor)
    multi = ParallelProcessor([])
    for frame_size in [1024, 2048, 4096]:
        frames = FramedSignalProcessor(frame_size=frame_size, fps=100)
        stft = ShortTimeFourierTransformProcessor()
        filt = FilteredSpectrogramProcessor(num_bands=6, fmin=30, fmax=17000, norm_filters=True)
        spec = LogarithmicSpectrogramProcessor(mul=mul, add=1)
        diff = SpectrogramDifferenceProcessor(diff_ratio=ratio, positive_diffs=True, stack_diffs=np.hstack)
        multi.append(SequentialProcessor((frames, stft, filt, spec, diff)))
    return SequentialProcessor((multi, np.hstack))


def test_multiresolution_stack(sig, ref_sample):
    """RNNOnsetProcessor / RNNBeatProcessor pre-processing spelled exactly as the reference does
    (features/onsets.py:778-791, features/beats.py:93-106) and through MultiFrameProcessor."""
    from madmom_b200.audio.spectrogram import MultiFrameProcessor
    on = _multires_chain(5, 0.25)(sig)
    assert on.shape == (281, 266)
    assert_close(on, ref_sample["rnn_onset_stack"], what="rnn onset stack")
    be = _multires_chain(1, 0.5)(sig)
    assert_close(be, ref_sample["rnn_beat_stack"], what="rnn beat stack")
    mf = MultiFrameProcessor(mul=5, diff_ratio=0.25)(sig)
    assert mf.shape == (281, 266) and mf.tensor.is_cuda
    assert np.array_equal(np.asarray(mf), on)


def test_onset_functions_and_peaks(sig, sample_pcm, ref_sample):
    from madmom_b200.audio.filters import LogarithmicFilterbank
    from madmom_b200.audio.signal import FramedSignal
    from madmom_b200.audio.spectrogram import Spectrogram, LogarithmicFilteredSpectrogram
    from madmom_b200.features.onsets import (spectral_flux, superflux, SpectralOnsetProcessor, peak_picking,
                                             OnsetPeakPickingProcessor)
    spec = Spectrogram(FramedSignal(sig, frame_size=2048, hop_size=441.))
    # spectral flux of the plain spectrogram (1024 columns: materialise + mmb_matrix_diff)
    sf = spectral_flux(spec)
    assert sf.shape == (281,)
    assert_close(np.asarray(sf), ref_sample["odf_spectralflux_plain"], what="spectral_flux(plain)")
    assert np.allclose(sf[:4], [0., 100.90121, 74.4442, 40.277565], rtol=1e-4)     # features/onsets.py:636-638
    lf24 = LogarithmicFilteredSpectrogram(spec, num_bands=24, norm_filters=False)
    assert_close(np.asarray(superflux(lf24)), ref_sample["odf_superflux_100"], what="superflux")
    lf12 = LogarithmicFilteredSpectrogram(spec, num_bands=12, norm_filters=False)
    assert_close(np.asarray(spectral_flux(lf12)), ref_sample["odf_logfiltspecflux"], what="log_filt_spec_flux")
    # bin/SuperFlux: fps 200 -> fractional hop 220.5, d = 2; onsets bit-exact
    proc = SpectralOnsetProcessor(onset_method="superflux", fps=200, filterbank=LogarithmicFilterbank, num_bands=24,
                                  fmin=30, fmax=17000, norm_filters=False, log=np.log10, mul=1, add=1,
                                  diff_ratio=0.5, diff_max_bins=3, positive_diffs=True)
    act = proc(sig)
    assert act.shape == (561,)
    assert_close(np.asarray(act), ref_sample["odf_superflux_200"], what="SpectralOnsetProcessor superflux")
    assert np.allclose(np.asarray(act), ref_sample["fixture_act_super_flux"], atol=1e-5)
    peaks = peak_picking(act, 1.1, None, 30, 0, 2, 10)
    assert np.array_equal(peaks, ref_sample["peaks_superflux_200"])
    pp = OnsetPeakPickingProcessor(threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                                   combine=0.03, delay=0, fps=200)
    onsets = pp(act)
    assert np.array_equal(onsets, ref_sample["onsets_superflux_200"])
    assert np.allclose(onsets, ref_sample["fixture_det_super_flux"], atol=1e-5)
    # bin/LogFiltSpecFlux detections
    pp = OnsetPeakPickingProcessor(threshold=1.6, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                                   combine=0.03, delay=0, fps=100)
    assert np.array_equal(pp(spectral_flux(lf12)), ref_sample["onsets_logfiltspecflux"])


def test_peak_picking_vs_oracle(ref_sample):
    """Integer output, bit-exact: literals of tests/test_features_onsets.py:233-245 and a parameter sweep against
    the scipy-based oracle on the reference's fixture and on random activations."""
    from oracle import frontend as O
    from madmom_b200 import _lib
    from madmom_b200.features.onsets import peak_picking
    act = ref_sample["fixture_act_super_flux"]
    onsets = peak_picking(act, 1.1)
    assert list(onsets[:6]) == [2, 10, 17, 48, 55, 80] and len(onsets) == 35
    assert list(peak_picking(act, 1.1, smooth=3)[:6]) == [2, 10, 17, 24, 48, 55]
    assert list(peak_picking(act, 1.1, pre_max=2, post_max=10, pre_avg=30)[:6]) == [2, 17, 55, 89, 122, 159]
    rng = np.random.default_rng(3)
    noise = (rng.random(4000).astype(np.float32) ** 4) * 3
    for a in (act, noise):
        for thr in (0.1, 0.5, 1.1):
            for smooth in (0, 3, 8):
                for pre_avg, post_avg in ((0, 0), (30, 0), (5, 7), (0, 9)):
                    for pre_max, post_max in ((1, 1), (0, 0), (2, 10), (7, 0)):
                        want = O.peak_picking(a, thr, smooth or None, pre_avg, post_avg, pre_max, post_max)
                        got = peak_picking(a, thr, smooth, pre_avg, post_avg, pre_max, post_max)
                        assert np.array_equal(got, want), (thr, smooth, pre_avg, post_avg, pre_max, post_max)
    # many activation functions in one call (segments), empty segment included
    lens = [561, 0, 4000, 17, 1]
    parts = [act, act[:0], noise, noise[100:117], noise[:1]]
    seg = torch.tensor(np.concatenate(([0], np.cumsum(lens))), dtype=torch.int64, device="cuda")
    peaks, counts = _lib.peak_picking(torch.from_numpy(np.concatenate(parts)).cuda(), 0.5, 0, 30, 0, 2, 10, seg)
    peaks, counts = peaks.cpu().numpy(), counts.cpu().numpy()
    for i, part in enumerate(parts):
        want = O.peak_picking(part, 0.5, None, 30, 0, 2, 10) if len(part) else np.zeros(0, int)
        s0 = int(seg[i])
        assert counts[i] == len(want) and np.array_equal(peaks[s0:s0 + counts[i]], want)


def test_frame_slices_and_origins(sig, sample_pcm):
    """FramedSignal.__getitem__(slice) (signal.py:1181-1194) and literal origins, fractional hop included: the
    GPU rows equal the oracle's on the same sliced framing."""
    from oracle import frontend as O
    from madmom_b200.audio.signal import FramedSignal
    from madmom_b200.audio.spectrogram import Spectrogram
    for hop in (441., 220.5):
        frames = FramedSignal(sig, frame_size=2048, hop_size=hop)
        whole = np.asarray(Spectrogram(frames))
        part = frames[10:50]
        assert part.num_frames == 40 and part.origin == int(0 - hop * 10)
        got = np.asarray(Spectrogram(part))
        fr = O.framed(sample_pcm, 2048, hop, origin=part.origin, n_frames=40)
        ref = O.spec(O.stft(fr, O.fft_window(np.hanning, 2048, np.int16)))
        assert np.abs(got - ref).max() <= 2e-6 * np.abs(ref).max()
        if hop == 441.:
            assert np.array_equal(got, whole[10:50])
    for origin in ("online", "stream"):
        frames = FramedSignal(sig, frame_size=2048, hop_size=441., origin=origin)
        got = np.asarray(Spectrogram(frames))
        fr = O.framed(sample_pcm, 2048, 441., origin=origin)
        ref = O.spec(O.stft(fr, O.fft_window(np.hanning, 2048, np.int16)))
        assert np.abs(got - ref).max() <= 2e-6 * np.abs(ref).max()


def test_float_signal_window_none(sample_pcm):
    from oracle import frontend as O
    from madmom_b200.audio.signal import Signal, FramedSignal
    from madmom_b200.audio.stft import ShortTimeFourierTransform
    from madmom_b200.audio.spectrogram import Spectrogram
    x = (sample_pcm.astype(np.float64) / 32767.).astype(np.float32)
    frames = FramedSignal(Signal(x, sample_rate=44100), frame_size=1024, hop_size=441.)
    got = np.asarray(Spectrogram(ShortTimeFourierTransform(frames, window=None)))
    ref = O.spec(O.stft(O.framed(x, 1024, 441.), None))
    assert np.abs(got - ref).max() <= 2e-6 * np.abs(ref).max()


def test_stft_attributes_and_scaled_signal(sig, sample_pcm):
    """The reference's own assertions on ShortTimeFourierTransform (tests/test_audio_stft.py:168-217): attribute
    values, STFT of an STFT, an int16 signal against the same signal scaled to -1..1 (the window takes the 1 / 32767),
    window=None, include_nyquist."""
    from madmom_b200.audio.signal import Signal
    from madmom_b200.audio.stft import ShortTimeFourierTransform, fft_frequencies
    from madmom_b200.audio.spectrogram import Spectrogram
    from madmom_b200.audio.stft import Phase
    result = ShortTimeFourierTransform(sig)
    assert result.shape == (281, 1024) and result.fft_size == 2048
    assert result.circular_shift is False and result.include_nyquist is False
    assert np.allclose(result.window, np.hanning(2048))
    assert np.allclose(result.fft_window, np.hanning(2048) / 32767)
    assert np.allclose(result.bin_frequencies, fft_frequencies(1024, 44100))
    assert result.num_frames == 281 and result.num_bins == 1024
    assert np.allclose(np.asarray(ShortTimeFourierTransform(result)), np.asarray(result))     # from STFT
    assert isinstance(result.spec(), Spectrogram) and isinstance(result.phase(), Phase)
    # test_fft_window: both STFTs must be the same
    scaling = float(np.iinfo(sample_pcm.dtype).max)
    scaled = Signal((sample_pcm / scaling).astype(np.float32), sample_rate=44100)
    scaled_result = ShortTimeFourierTransform(scaled)
    assert np.allclose(np.asarray(result), np.asarray(scaled_result), atol=1e-6)
    plain = ShortTimeFourierTransform(sig, window=None)
    assert np.allclose(plain.fft_window, np.ones(2048, dtype=float) / scaling)
    assert ShortTimeFourierTransform(scaled, window=None).fft_window is None
    assert np.allclose(np.asarray(plain), np.asarray(ShortTimeFourierTransform(scaled, window=None)), atol=2e-5)
    # test_nyquist
    nyq = ShortTimeFourierTransform(sig, include_nyquist=True)
    assert nyq.shape == (281, 1025) and nyq.fft_size == 2048 and nyq.include_nyquist is True
    assert np.allclose(nyq.bin_frequencies, fft_frequencies(1025, 44100))
    assert np.array_equal(np.asarray(nyq)[:, :1024], np.asarray(result))


def test_mel_filterbank_and_cnn_front_end(sig, ref_mel):
    """SURVEY 8(f) row 2 against vectors of the real reference (tests/golden/make_golden_mel.py):
    FilteredSpectrogram(MelFilterbank, 40) with the literals of tests/test_audio_spectrogram.py:173-180, a
    rectangular (Bark) bank as a non-triangular case, and the CNNOnsetProcessor front end spelled as the reference
    does (features/onsets.py:838-871): Mel 80 bands 27.5 Hz .. 16 kHz with duplicate low bands, ln(x + EPSILON),
    np.dstack over frame sizes [2048, 1024, 4096].  Standard bound, no mask."""
    from madmom_b200.processors import ParallelProcessor, SequentialProcessor
    from madmom_b200.audio.filters import Filterbank, MelFilterbank
    from madmom_b200.audio.signal import FramedSignalProcessor, SignalProcessor
    from madmom_b200.audio.stft import ShortTimeFourierTransformProcessor
    from madmom_b200.audio.spectrogram import (FilteredSpectrogram, FilteredSpectrogramProcessor,
                                               LogarithmicSpectrogramProcessor)
    from madmom_b200.recipe import fft_frequencies
    mel = FilteredSpectrogram(sig, filterbank=MelFilterbank, num_bands=40)
    assert mel.shape == (281, 40)
    assert np.allclose(mel[0, :6], [8.42887115, 17.98174477, 19.50165367, 6.48194313, 2.96991181, 4.06280804])
    assert_close(np.asarray(mel), ref_mel["filt_mel40"], what="FilteredSpectrogram(MelFilterbank)")
    bark = Filterbank(ref_mel["bark_N2048_dense"], fft_frequencies(1024, 44100))
    got = FilteredSpectrogram(sig, filterbank=bark)
    assert np.allclose(got[0, :6], [16.42251968, 17.36715126, 2.81979132, 4.27050114, 3.08699131, 1.50553513])
    assert_close(np.asarray(got), ref_mel["filt_bark"], what="FilteredSpectrogram(rectangular bank)")
    EPSILON = np.spacing(1)
    multi = ParallelProcessor([])
    for frame_size in [2048, 1024, 4096]:
        frames = FramedSignalProcessor(frame_size=frame_size, fps=100)
        stft = ShortTimeFourierTransformProcessor()
        filt = FilteredSpectrogramProcessor(filterbank=MelFilterbank, num_bands=80, fmin=27.5, fmax=16000,
                                            norm_filters=True, unique_filters=False)
        spec = LogarithmicSpectrogramProcessor(log=np.log, add=EPSILON)
        multi.append(SequentialProcessor((frames, stft, filt, spec)))
    pre = SequentialProcessor((SignalProcessor(num_channels=1, sample_rate=44100), multi, np.dstack))
    out = pre(sig)
    assert out.shape == (281, 80, 3) and out.dtype == np.float32
    assert_close(out, ref_mel["cnn_onset_frontend"], what="CNN onset front end")


def test_spectrogram_of_a_spectrogram_keeps_the_data(sig, ref_sample):
    """Spectrogram(x) / SpectrogramProcessor()(x) for x already a (filtered / logarithmic) Spectrogram keeps x's
    matrix and only re-types it (madmom/audio/spectrogram.py:125-127)."""
    from madmom_b200.audio.spectrogram import (Spectrogram, SpectrogramProcessor, LogarithmicFilteredSpectrogram,
                                               FilteredSpectrogram, LogarithmicSpectrogram)
    lfs = LogarithmicFilteredSpectrogram(sig)
    again = SpectrogramProcessor()(lfs)
    assert type(again) is Spectrogram and again.shape == (281, 81) and again.num_bins == 81
    assert_close(np.asarray(again), ref_sample["logfilt12n"], what="Spectrogram(LogarithmicFilteredSpectrogram)")
    assert again.stft is lfs.stft
    # the re-typed matrix keeps going like the reference's: its `bin_frequencies` are the STFT's 1024 (a base
    # Spectrogram has no other, spectrogram.py:143-147), so a new filterbank does not fit the 81 columns (np.dot
    # raises ValueError there too); the logarithm of the logarithm is just computed (tests/test_gpu_reference_suite.py)
    with pytest.raises(ValueError):
        FilteredSpectrogram(again)
    assert np.allclose(np.asarray(LogarithmicSpectrogram(again)), np.log10(np.asarray(again) + 1), rtol=1e-5,
                       atol=1e-6)


def test_errors(sig):
    from madmom_b200.audio.signal import Signal, FramedSignal
    from madmom_b200.audio.stft import ShortTimeFourierTransform
    from madmom_b200.audio.spectrogram import FilteredSpectrogram, SpectrogramDifferenceProcessor, Spectrogram
    stereo = Signal(np.zeros((5000, 2), np.int16), sample_rate=44100)
    with pytest.raises(ValueError):
        ShortTimeFourierTransform(FramedSignal(stereo))              # stft.py:81-84
    with pytest.raises(TypeError):
        FilteredSpectrogram(FramedSignal(sig), filterbank="nope")    # spectrogram.py:378-380
    with pytest.raises(ValueError):
        FramedSignal(sig, end="bogus")                               # signal.py:1153-1155
    with pytest.raises(NotImplementedError):
        ShortTimeFourierTransform(FramedSignal(sig, frame_size=512))
    with pytest.raises(NotImplementedError):
        ShortTimeFourierTransform(FramedSignal(sig), fft_size=4096)
    with pytest.raises(NotImplementedError):
        SpectrogramDifferenceProcessor()(Spectrogram(FramedSignal(sig)).filter(), reset=False)
    # mono down-mix like the reference (mean cast back to int16) then the GPU path
    mixed = Signal(np.stack((np.asarray(sig), np.asarray(sig) // 2), axis=1), sample_rate=44100, num_channels=1)
    assert mixed.ndim == 1 and Spectrogram(FramedSignal(mixed)).shape == (281, 1024)
    before = np.asarray(sig).copy()
    np.asarray(Spectrogram(FramedSignal(sig)))
    assert np.array_equal(before, np.asarray(sig))                   # inputs are never mutated


def test_long_signal_shards_on_one_gpu():
    """The per-rank work of the long-signal path (parallel.LongSignalFrontEnd.run_local) for 1, 2, 4 and 8 ranks,
    executed one rank after the other on this GPU and concatenated like the all-gather does, is bit-identical to
    the unsharded launch (fractional hop and d = 2 included)."""
    from madmom_b200 import parallel, recipe as R
    from madmom_b200.batch import BatchFrontEnd
    from madmom_b200.synth import make_track
    x = make_track(11, num_samples=400003)
    for rec in (R.SUPERFLUX, R.SUPERFLUX_200) + tuple(R.RNN_ONSET):       # RNN_ONSET: N = 1024 / 2048 / 4096
        whole = BatchFrontEnd(rec).run_device(torch.from_numpy(x).cuda(), [len(x)], ("diff", "odf"))
        T = whole["diff"].shape[0]
        fe = parallel.LongSignalFrontEnd(rec)
        for world in (1, 2, 4, 8):
            parts = {"diff": [], "odf": []}
            for rank in range(world):
                sh = fe.shard(len(x), world, rank)
                local = fe.run_local(torch.from_numpy(x[sh.sample_lo:sh.sample_hi].copy()).cuda(), sh,
                                     ("diff", "odf"))
                for k in parts:
                    parts[k].append(local[k])
            for k in parts:
                full = torch.cat(parts[k])[:T]
                assert torch.equal(full, whole[k]), (rec.hop_size, world, k)


def test_run_gather_stores_every_row_to_all_destinations():
    """mmb_run_gather (the front end fused with the all-gather of a sharded long signal): with three destinations
    on this GPU standing in for the ranks' buffers, every destination receives the rows mmb_run produces, bit for
    bit, for each shard of a 4-rank split; misuse is rejected."""
    from madmom_b200 import _lib, parallel, recipe as R
    from madmom_b200.synth import make_track
    x = make_track(12, num_samples=300001)
    for rec in (R.SUPERFLUX_200, R.RNN_ONSET[0]):
        fe = parallel.LongSignalFrontEnd(rec)
        B = fe.plan.num_bands
        for rank in range(4):
            sh = fe.shard(len(x), 4, rank)
            pcm = torch.from_numpy(x[sh.sample_lo:sh.sample_hi].copy()).cuda()
            want = fe.run_local(pcm, sh, ("log", "diff", "odf"))
            rows = sh.rows_per_rank
            dst = [{"log": torch.zeros((rows, B), device="cuda"), "diff": torch.zeros((rows, B), device="cuda"),
                    "odf": torch.zeros((rows,), device="cuda")} for _ in range(3)]
            units = np.zeros(1, dtype=_lib.TRACK_DTYPE)
            units[0] = sh.unit()
            n = fe.plan.run_gather(pcm, units, [{k: (t.data_ptr(), t.stride(0) if t.dim() == 2 else 1)
                                                 for k, t in d.items()} for d in dst])
            assert n == 1
            for d in dst:
                for k in ("log", "diff", "odf"):
                    assert torch.equal(d[k], want[k]), (rec.frame_size, rank, k)
    with pytest.raises(ValueError):
        fe.plan.run_gather(pcm, units, [{"spec": (dst[0]["log"].data_ptr(), B)}])
    with pytest.raises((ValueError, _lib.LibraryError)):          # destinations must agree on the leading dimension
        fe.plan.run_gather(pcm, units, [{"diff": (dst[0]["diff"].data_ptr(), B)},
                                        {"diff": (dst[1]["diff"].data_ptr(), B + 1)}])


def test_batch_front_end_stack_and_host_path():
    from oracle import frontend as O
    from madmom_b200 import recipe as R
    from madmom_b200.batch import BatchFrontEnd
    from madmom_b200.synth import make_track
    tracks = [make_track(1000 + i, num_samples=n) for i, n in enumerate((30000, 44100, 12345))]
    pcm = np.concatenate(tracks)
    lens = [len(t) for t in tracks]
    fe = BatchFrontEnd(R.RNN_BEAT)
    out = fe.run_device(torch.from_numpy(pcm).cuda(), lens, ("stack",))["stack"].cpu().numpy()
    assert out.shape[1] == 266
    r0 = 0
    for x in tracks:
        cols = []
        for N in (1024, 2048, 4096):
            st = O.run_recipe(x, dict(frame_size=N, hop_size=441., num_bands=6, norm_filters=True, mul=1, add=1,
                                      diff_ratio=0.5, positive_diffs=True), stages=True)
            cols += [st["log"], st["diff"]]
        want = np.hstack(cols)
        assert_close(out[r0:r0 + len(want)], want, what="batched multi-resolution stack")
        r0 += len(want)
    assert r0 == out.shape[0]
    host = fe.run_host(pcm, lens, ("stack",))["stack"].numpy()
    assert np.array_equal(host, out)


def test_complex_flux(sig, ref_sample):
    """bin/ComplexFlux recipe (fps 200, circular shift, d = 2): local group delay rows against the reference's
    unwrap-based values, ODF within tolerance of the reference and of its shipped fixture, onsets bit-exact.
    The reference's lgd carries the rounding of a float32 cumulative unwrap (up to 4e-5 on |lgd|/pi); the kernel's
    values are the exact principal differences, hence the looser bound on the raw lgd rows."""
    from madmom_b200.audio.filters import LogarithmicFilterbank
    from madmom_b200.audio.signal import FramedSignal
    from madmom_b200.audio.stft import ShortTimeFourierTransform
    from madmom_b200.audio.spectrogram import Spectrogram, LogarithmicFilteredSpectrogram
    from madmom_b200.features.onsets import complex_flux, SpectralOnsetProcessor, OnsetPeakPickingProcessor
    kw = dict(fps=200, filterbank=LogarithmicFilterbank, num_bands=24, fmin=30, fmax=17000, norm_filters=False,
              log=np.log10, mul=1, add=1, diff_ratio=0.5, diff_max_bins=3, positive_diffs=True)
    act = SpectralOnsetProcessor(onset_method="complex_flux", **kw)(sig)
    assert act.shape == (561,)
    assert_close(np.asarray(act), ref_sample["odf_complexflux_200"], what="complex_flux")
    assert np.allclose(np.asarray(act), ref_sample["fixture_act_complex_flux"], atol=2e-5)    # tests/test_bin.py:350-386
    pp = OnsetPeakPickingProcessor(threshold=0.25, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                                   combine=0.03, delay=0, fps=200)
    onsets = pp(act)
    assert np.array_equal(onsets, ref_sample["onsets_complexflux_200"])
    assert np.allclose(onsets, ref_sample["fixture_det_complex_flux"], atol=1e-5)
    # the lgd rows themselves
    frames = FramedSignal(sig, frame_size=2048, fps=200)
    spec = LogarithmicFilteredSpectrogram(Spectrogram(ShortTimeFourierTransform(frames, circular_shift=True)),
                                          num_bands=24, norm_filters=False)
    cf = complex_flux(spec)
    lgd = cf._chain.run(("lgd",))["lgd"].cpu().numpy()
    want = np.abs(ref_sample["lgd_200_rows"]) / np.pi
    err = np.abs(lgd[ref_sample["rows"]] - want)
    # bins above ~13 kHz of sample.wav are 70 dB down: their phase is rounding noise of the FP32 / float64 FFTs
    assert err[:, :600].max() < 1e-4 and np.median(err) < 1e-6 and err.max() < 1e-2
    assert not lgd[:, -1].any()
    # tests/test_features_onsets.py:62-86 literal (norm_filters=True, fps 100)
    spec = LogarithmicFilteredSpectrogram(Spectrogram(FramedSignal(sig), circular_shift=True), num_bands=24)
    assert np.allclose(complex_flux(spec)[:6], [0, 0.476213485, 0.0877621323, 0.0593151376, 0.0654867291,
                                                0.0954693183], atol=2e-5)


def test_phase_and_local_group_delay(sig, ref_phase):
    """Phase / LocalGroupDelay (madmom/audio/stft.py:556-708) against rows computed by the real reference
    (tests/golden/make_golden_phase.py): the phase to 2e-6 rad modulo 2 pi (the bins come from the float64 kernel),
    the local group delay to 1e-4 (the reference's float32 cumulative unwrap carries up to about 4e-5 of rounding;
    here it is the wrapped phase difference itself).  circular_shift defaults to True when the STFT is built by
    Phase; an STFT without it draws the reference's warning; chaining and aliases as in the reference."""
    import warnings
    from madmom_b200.audio.stft import (ShortTimeFourierTransform, Phase, LocalGroupDelay, LGD, phase,
                                        local_group_delay)
    rows = ref_phase["rows"]

    def wrapped(a, b):
        d = np.abs(a.astype(np.float64) - b.astype(np.float64))
        return np.minimum(d, 2 * np.pi - d)
    ph = Phase(sig)
    assert ph.stft.circular_shift and ph.shape == (281, 1024) and ph.dtype == np.float32
    assert wrapped(np.asarray(ph)[rows], ref_phase["phase_rows"]).max() <= 2e-6
    lg = ph.lgd()
    assert isinstance(lg, LocalGroupDelay) and LGD is LocalGroupDelay and lg.stft is ph.stft and lg.phase is ph
    assert lg.shape == (281, 1024) and not np.asarray(lg)[:, -1].any()
    assert wrapped(np.asarray(lg)[rows], ref_phase["lgd_rows"]).max() <= 1e-4
    assert wrapped(np.asarray(LocalGroupDelay(sig))[rows], ref_phase["lgd_rows"]).max() <= 1e-4
    assert np.allclose(ph.bin_frequencies, ph.stft.bin_frequencies)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        plain = Phase(ShortTimeFourierTransform(sig))                  # no circular shift: the docstring's example
        assert any(issubclass(x.category, RuntimeWarning) for x in w)
    assert wrapped(np.asarray(plain)[rows], ref_phase["phase_plain_rows"]).max() <= 2e-6
    # bin 0 is real: +pi for a negative sum like np.angle gives, not -pi
    for got, want in ((np.asarray(plain)[rows], ref_phase["phase_plain_rows"]),
                      (np.asarray(ph)[rows], ref_phase["phase_rows"])):
        assert np.array_equal(got[:, 0], want[:, 0])
    # function forms and the method on the STFT
    assert np.array_equal(np.asarray(phase(ph.stft)), np.asarray(ph))
    assert np.array_equal(np.asarray(local_group_delay(ph)), np.asarray(lg))
    assert np.array_equal(np.asarray(ph.stft.phase()), np.asarray(ph))
    with pytest.raises(ValueError):
        local_group_delay(np.arange(10))


@pytest.mark.parametrize("which", ["sample", "synth"])
def test_other_spectral_odfs(sample_pcm, ref_odfs, which):
    """the remaining onset detection functions of madmom/features/onsets.py, with the inputs the
    reference's tests use (tests/test_features_onsets.py:24-27), against vectors computed by the real reference.
    The phase-based functions average over all 1024 bins, noise-floor bins included, so they need the raw bins as the
    reference has them: the stft / spec matrices come from the float64 kernel (csrc/stft64.cu), and every function
    meets the standard bound 1e-5 + 1e-4 |ref| on both inputs."""
    from madmom_b200.audio.signal import Signal
    from madmom_b200.audio.spectrogram import Spectrogram, LogarithmicFilteredSpectrogram
    from madmom_b200.features import onsets as ON
    from madmom_b200.synth import make_track
    x = sample_pcm if which == "sample" else make_track(3, 2.0)
    spec = Spectrogram(Signal(x, sample_rate=44100), circular_shift=True)
    lfs = LogarithmicFilteredSpectrogram(spec, num_bands=24, mul=1, add=1)
    got = {"hfc_logfilt": ON.high_frequency_content(lfs), "hfc_spec": ON.high_frequency_content(spec),
           "sd_logfilt": ON.spectral_diff(lfs), "sd_spec": ON.spectral_diff(spec),
           "mkl_logfilt": ON.modified_kullback_leibler(lfs), "mkl_spec": ON.modified_kullback_leibler(spec),
           "pd": ON.phase_deviation(lfs), "wpd": ON.weighted_phase_deviation(spec),
           "nwpd": ON.normalized_weighted_phase_deviation(spec), "cd": ON.complex_domain(spec),
           "rcd": ON.rectified_complex_domain(spec)}
    worst = {}
    for k, v in got.items():
        want = ref_odfs["%s_%s" % (which, k)]
        a = np.asarray(v)
        assert a.shape == want.shape, k
        worst[k] = float(np.max(np.abs(a - want) / (1e-5 + 1e-4 * np.abs(want))))
    print(which, {k: round(v, 3) for k, v in worst.items()})
    for k in worst:
        assert worst[k] <= 1.0, (k, worst[k])
    if which == "sample":
        # reference literals, tests/test_features_onsets.py:35-152
        assert np.allclose(np.asarray(got["hfc_logfilt"])[:6], [8.97001563, 9.36399107, 8.64144536, 8.34977449,
                                                               8.21097918, 8.40412515], rtol=1e-4)
        assert np.allclose(np.asarray(got["cd"])[:6], [399.29980469, 585.9564209, 262.08010864, 225.84718323,
                                                      196.88954163, 200.32469177], rtol=1e-4)
        assert np.allclose(np.asarray(got["wpd"])[:6], [0, 0, 0.19568817, 0.20483065, 0.17890805, 0.16970603],
                           rtol=1e-4)


def test_spectral_odf_errors(sig):
    from madmom_b200.audio.spectrogram import Spectrogram, LogarithmicFilteredSpectrogram
    from madmom_b200.features import onsets as ON
    lfs = LogarithmicFilteredSpectrogram(Spectrogram(sig, circular_shift=True), num_bands=24)
    # reference: tests/test_features_onsets.py:108-152
    for fn in (ON.weighted_phase_deviation, ON.normalized_weighted_phase_deviation, ON.complex_domain,
               ON.rectified_complex_domain):
        with pytest.raises(ValueError):
            fn(lfs)
    with pytest.raises(ValueError):
        ON.modified_kullback_leibler(lfs, epsilon=0)


def test_pcm_ingest_remix():
    """interleaved multi-channel PCM mixed down on the device, bit for bit like
    madmom.audio.signal.remix (np.mean(signal, axis=-1).astype(dtype), signal.py:212; its known answers
    tests/test_audio_signal.py:277-361 are integer truncation cases like the ones below)."""
    from madmom_b200 import _lib, recipe as R
    from madmom_b200.batch import BatchFrontEnd
    rng = np.random.default_rng(5)
    st = rng.integers(-32768, 32768, size=(50001, 2)).astype(np.int16)
    st[:4] = [[1, 2], [-1, -2], [32767, 32767], [-32768, -32767]]            # 1.5 -> 1, -1.5 -> -1, no clipping
    want = np.mean(st, axis=-1).astype(np.int16)
    got = _lib.remix_mono(torch.from_numpy(st).cuda()).cpu().numpy()
    assert np.array_equal(got, want) and list(got[:4]) == [1, -1, 32767, -32767]
    assert np.array_equal(_lib.remix_mono(torch.from_numpy(st).cuda(), channel=1).cpu().numpy(), st[:, 1])
    six = rng.integers(-32768, 32768, size=(7777, 6)).astype(np.int16)
    assert np.array_equal(_lib.remix_mono(torch.from_numpy(six).cuda()).cpu().numpy(),
                          np.mean(six, axis=-1).astype(np.int16))
    fl = rng.standard_normal((40000, 3)).astype(np.float32)
    assert np.array_equal(_lib.remix_mono(torch.from_numpy(fl).cuda()).cpu().numpy(), np.mean(fl, axis=-1))
    # host pipeline: interleaved stereo in, features of the mono mix out
    fe = BatchFrontEnd(R.SUPERFLUX)
    lens = [30000, 20001]
    stereo = st[:sum(lens)]
    mono = np.mean(stereo, axis=-1).astype(np.int16)
    a = fe.run_host(stereo, lens, ("diff",), num_channels=2)["diff"].numpy()
    b = fe.run_host(mono, lens, ("diff",))["diff"].numpy()
    assert a.shape == b.shape and np.array_equal(a, b)
    # a second call reuses the plan's pipeline buffers
    assert np.array_equal(fe.run_host(stereo, lens, ("diff",), num_channels=2)["diff"].numpy(), b)


@pytest.mark.gpu
@pytest.mark.parametrize("n_bins,B", [(1024, 77), (512, 77), (640, 77), (2048, 77), (512, 100), (1024, 128), (1024, 140),
                                      (2048, 200)])
def test_complex_flux_kernel_ragged_signals(n_bins, B, monkeypatch):
    """mmb_complex_flux on random matrices with signals of 1, 2, 3 and many rows back to back (temporal 'reflect'
    maximum inside each signal, madmom/features/onsets.py:282-317): the block kernel (one CTA per run of rows,
    hierarchical band minima; four, two or one lane per band depending on bins and bands), the streaming kernel
    (one warp per row) and the generic kernel give the same bits, and match scipy's maximum_filter + the per-band
    minimum.  Band 7 is wide enough to overflow the block kernel's register cover (its scan path)."""
    torch = pytest.importorskip("torch")
    from scipy.ndimage import maximum_filter
    from madmom_b200 import _lib
    rng = np.random.default_rng(n_bins + B)
    lens = [1, 2, 3, 1, 57, 700, 2, 333, 1]
    rows = sum(lens)
    lgd = rng.random((rows, n_bins), dtype=np.float32)
    diff = rng.random((rows, B), dtype=np.float32)
    centres = np.unique(np.round(np.geomspace(2, n_bins - 3, B + 2)).astype(np.int64))
    lo = np.zeros(B, np.int32); hi = np.zeros(B, np.int32)
    for b in range(B):
        c0 = centres[min(b, len(centres) - 3)]; c2 = centres[min(b, len(centres) - 3) + 2]
        lo[b], hi[b] = max(c0 - 1, 0), min(c2 + 2, n_bins)
    hi[5] = lo[5]                                                    # an empty band contributes nothing
    lo[7], hi[7] = 3, n_bins - 5                                     # a cover of more than 16 / 32 blocks
    seg = np.concatenate(([0], np.cumsum(lens))).astype(np.int64)
    want = np.zeros(rows, np.float64)
    for s in range(len(lens)):
        a, e = seg[s], seg[s + 1]
        mx = maximum_filter(lgd[a:e], size=[3, 1], origin=0)
        mask = np.stack([mx[:, lo[b]:hi[b]].min(axis=1) if hi[b] > lo[b] else np.zeros(e - a, np.float32)
                         for b in range(B)], axis=1)
        want[a:e] = (diff[a:e].astype(np.float64) * mask).sum(axis=1)
    dev = torch.device("cuda:0")
    t = lambda x: torch.from_numpy(x).to(dev)
    args = (t(lgd), t(diff), t(lo), t(hi))

    def other(env, **kw):
        monkeypatch.setenv(env, "1")
        try:
            return _lib.complex_flux(*kw.pop("args", args), **kw).cpu().numpy()
        finally:
            monkeypatch.delenv(env)

    got = _lib.complex_flux(*args, seg_offsets=t(seg)).cpu().numpy()
    assert np.allclose(got, want, rtol=1e-5, atol=1e-5)
    assert np.array_equal(got, other("MMB_CF_GENERIC", seg_offsets=t(seg)))
    assert np.array_equal(got, other("MMB_CF_STREAM", seg_offsets=t(seg)))      # generic again for 640 / 2048 bins
    # equal-length signals by rows_per_signal, and one signal only
    args2 = (t(lgd[:1000]), t(diff[:1000]), args[2], args[3])
    for rps in (0, 8, 250, 300, 999):                 # 300, 999: the last signal is shorter than rows_per_signal
        a = _lib.complex_flux(*args2, rows_per_signal=rps).cpu().numpy()
        assert np.array_equal(a, other("MMB_CF_GENERIC", args=args2, rows_per_signal=rps)), rps
        assert np.array_equal(a, other("MMB_CF_STREAM", args=args2, rows_per_signal=rps)), rps
    # a run boundary in every position relative to a signal: 70 000 rows in signals of 37 (the block kernel's runs
    # are 64 rows or longer)
    big_l, big_d = t(np.tile(lgd, (64, 1))[:70000].copy()), t(np.tile(diff, (64, 1))[:70000].copy())
    args3 = (big_l, big_d, args[2], args[3])
    a = _lib.complex_flux(*args3, rows_per_signal=37).cpu().numpy()
    assert np.array_equal(a, other("MMB_CF_GENERIC", args=args3, rows_per_signal=37))


@pytest.mark.gpu
@pytest.mark.parametrize("cols", [1024, 512, 1025])
def test_phase_deviation_kernels_against_numpy(cols):
    """The three phase-deviation functions on random complex matrices with signals of 1, 2, 3 and many rows back to
    back against numpy's definition (madmom/features/onsets.py:320-437: mean |wrap(phi[t] - 2 phi[t-1] + phi[t-2])|,
    weighted by |X|, normalised), within 1e-5 + 1e-4 |ref| for every row.  The kernel takes the wrapped second
    difference as the argument of x0 conj(x1)^2 x2 (one arctangent per bin): bins of very different magnitude, a
    silent frame (phase 0 like np.angle(0)) and bins on the negative real axis are in the input."""
    from madmom_b200 import _lib
    rng = np.random.default_rng(cols)
    lens = [1, 2, 3, 1, 400, 2, 1700, 333, 1]
    rows = sum(lens)
    x = (rng.standard_normal((rows, cols)) + 1j * rng.standard_normal((rows, cols))).astype(np.complex64)
    x *= (10.0 ** rng.uniform(-12, 3, size=(rows, cols))).astype(np.float32)      # 300 dB of dynamic range
    x[5, :] = 0
    x[40, 3] = -1.0
    x[41:44, 7] = [-2.0, -3.0, -0.5]
    seg = np.concatenate(([0], np.cumsum(lens))).astype(np.int64)
    dev = torch.device("cuda:0")
    xs, segs = torch.from_numpy(x).to(dev), torch.from_numpy(seg).to(dev)
    for mode in ("phase_deviation", "weighted_phase_deviation", "normalized_weighted_phase_deviation"):
        got = _lib.spectral_odf(mode, stft=xs, seg_offsets=segs).cpu().numpy()
        want = np.zeros(rows)
        for s in range(len(lens)):
            a, e = seg[s], seg[s + 1]
            ph = np.angle(x[a:e].astype(np.complex128))
            pd = np.zeros_like(ph)
            if e - a > 2:
                pd[2:] = ph[2:] - 2 * ph[1:-1] + ph[:-2]
            pd = np.mod(pd + np.pi, 2 * np.pi) - np.pi
            mag = np.abs(x[a:e]).astype(np.float64)
            if mode == "phase_deviation":
                want[a:e] = np.mean(np.abs(pd), axis=1)
            elif mode == "weighted_phase_deviation":
                want[a:e] = np.mean(np.abs(pd * mag), axis=1)
            else:
                want[a:e] = np.mean(np.abs(pd * mag), axis=1) / (np.mean(mag, axis=1) + np.finfo(float).eps)
        err = np.abs(got - want) - (1e-5 + 1e-4 * np.abs(want))
        assert err.max() <= 0, (mode, int(err.argmax()), got[err.argmax()], want[err.argmax()])


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(int(__import__("os").environ.get("MMB_FUZZ_SEEDS", "24"))))
def test_random_long_signal_shards_are_bit_identical(seed):
    """The property the multi-GPU path rests on, over random cases: a signal of random length cut into the frame
    ranges of 2 .. 9 ranks (more ranks than frames included), any frame size, integer and fractional hops, lags of
    1 .. 3 frames, with and without the frequency max-filter -- each rank's rows computed from its own PCM slice with
    its halo (parallel.LongSignalFrontEnd.run_local), concatenated like the all-gather does -- equals the unsharded
    launch bit for bit, float32 PCM included."""
    from madmom_b200 import parallel, recipe as R
    from madmom_b200.batch import BatchFrontEnd
    from madmom_b200.synth import make_track
    rng = np.random.default_rng(8000 + seed)
    N = int(rng.choice([1024, 2048, 4096]))
    hop = float(rng.choice([441., 220.5, 882., 300.25, 147., 1000.]))
    rec = R.SUPERFLUX.replace(frame_size=N, hop_size=hop, num_bands=int(rng.choice([6, 12, 24])),
                              diff_ratio=float(rng.choice([0.25, 0.5, 0.75])),
                              diff_max_bins=[None, 3][int(rng.integers(2))])
    n = int(rng.choice([1, 17, int(hop), N - 1, N + 1, int(rng.integers(2000, 300000)), int(rng.integers(2000, 300000))]))
    dtype = np.float32 if rng.integers(4) == 0 else np.int16
    x = make_track(12000 + seed, num_samples=n)
    if dtype == np.float32:
        x = (x / 32767.).astype(np.float32)
    whole = BatchFrontEnd(rec, dtype).run_device(torch.from_numpy(x).cuda(), [len(x)], ("diff", "odf"))
    T = whole["diff"].shape[0]
    fe = parallel.LongSignalFrontEnd(rec, dtype)
    for world in sorted(set(int(w) for w in rng.integers(2, 10, size=3))):
        parts = {"diff": [], "odf": []}
        covered = 0
        for rank in range(world):
            sh = fe.shard(len(x), world, rank)
            assert sh.frame_first == min(covered, T) or sh.frame_count == 0
            covered += sh.frame_count
            local = fe.run_local(torch.from_numpy(x[sh.sample_lo:sh.sample_hi].copy()).cuda(), sh, ("diff", "odf"))
            for k in parts:
                parts[k].append(local[k])
        assert covered == T
        for k in parts:
            full = torch.cat(parts[k])[:T]
            assert torch.equal(full, whole[k]), "seed %d N=%d hop=%g n=%d world=%d %s %s" % (
                seed, N, hop, n, world, np.dtype(dtype).name, k)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(int(__import__("os").environ.get("MMB_FUZZ_SEEDS", "33"))))
def test_random_onset_processor_chains_vs_oracle(seed):
    """SpectralOnsetProcessor over its options -- every onset method, 100 / 200 fps, with and without a logarithmic
    filterbank (6 / 12 / 24 bands per octave, normalised or not), with and without logarithmic scaling -- on seeded
    synthetic signals of random length, against the oracle's composition of the same stages (the processor calls a
    method with its defaults, madmom/features/onsets.py:651-684).  Same bound as everywhere; 33 seeds in the suite
    (each method three times), 220 were run once in round 2."""
    from oracle import frontend as O
    from madmom_b200.audio.signal import Signal
    from madmom_b200.audio.filters import LogarithmicFilterbank
    from madmom_b200.features.onsets import SpectralOnsetProcessor
    from madmom_b200.synth import make_track
    methods = ["superflux", "complex_flux", "high_frequency_content", "spectral_diff", "spectral_flux",
               "modified_kullback_leibler", "phase_deviation", "weighted_phase_deviation",
               "normalized_weighted_phase_deviation", "complex_domain", "rectified_complex_domain"]
    rng = np.random.default_rng(11000 + seed)
    method = methods[seed % len(methods)]
    fps = int(rng.choice([100, 200]))
    unfiltered_only = method in ("weighted_phase_deviation", "normalized_weighted_phase_deviation", "complex_domain",
                                 "rectified_complex_domain")
    filtered = method == "complex_flux" or (not unfiltered_only and bool(rng.integers(3)))
    bands, norm = int(rng.choice([6, 12, 24])), bool(rng.integers(2))
    log = method == "complex_flux" or bool(rng.integers(2))
    mul = float(rng.choice([1., 5.]))
    x = make_track(14000 + seed, num_samples=int(rng.integers(30000, 140000)))
    kw = dict(onset_method=method, fps=fps)
    if filtered:
        kw.update(filterbank=LogarithmicFilterbank, num_bands=bands, norm_filters=norm)
    if log:
        kw.update(log=np.log10, mul=mul, add=1)
    got = np.asarray(SpectralOnsetProcessor(**kw)(Signal(x, sample_rate=44100)))
    # ---- the same chain on the oracle
    hop = 44100. / fps
    circ = method in ("phase_deviation", "weighted_phase_deviation", "normalized_weighted_phase_deviation",
                      "complex_domain", "rectified_complex_domain", "complex_flux")
    st = O.stft(O.framed(x, 2048, hop), O.fft_window(np.hanning, 2048, x.dtype), circular_shift=circ)
    sp = O.spec(st)
    fb = None
    if filtered:
        fb = O.logarithmic_filterbank(O.fft_frequencies(1024, 44100), bands, 30., 17000., 440., norm, True)
        sp = O.filtered(sp, fb)
    if log:
        sp = O.logarithmic(sp, mul, 1.)
    d = O.diff_frames(0.5, hop, 2048, np.hanning(2048))
    want = {"superflux": lambda: O.flux(O.difference(sp, d, 3, True)),
            "spectral_flux": lambda: O.flux(O.difference(sp, d, None, True)),
            "complex_flux": lambda: O.complex_flux(sp, st, fb, d),
            "high_frequency_content": lambda: O.high_frequency_content(sp),
            "spectral_diff": lambda: O.spectral_diff(sp, d),
            "modified_kullback_leibler": lambda: O.modified_kullback_leibler(sp, 1),
            "phase_deviation": lambda: O.phase_deviation(st),
            "weighted_phase_deviation": lambda: O.weighted_phase_deviation(st),
            "normalized_weighted_phase_deviation": lambda: O.normalized_weighted_phase_deviation(st),
            "complex_domain": lambda: O.complex_domain(st),
            "rectified_complex_domain": lambda: O.rectified_complex_domain(st, d)}[method]()
    assert got.shape == np.shape(want)
    # Without a logarithm the functions that reduce the spectrogram add up values of the order of its largest entry
    # (hundreds for int16 input), each with a float32 rounding of that size -- in the reference as much as here -- so
    # the absolute term of the bound is taken relative to that scale (squared for spectral_diff, which squares).
    # Seed 132 of a 220-seed run (superflux of a filtered, unscaled spectrogram) is the case that needs it.
    scale = 1.0
    if not log and method in ("superflux", "spectral_flux", "spectral_diff", "high_frequency_content"):
        scale = max(1.0, float(np.abs(sp).max())) ** (2 if method == "spectral_diff" else 1)
    what = "seed %d %s fps=%d filtered=%s bands=%d norm=%s log=%s mul=%g n=%d" % (
        seed, method, fps, filtered, bands, norm, log, mul, len(x))
    if method == "rectified_complex_domain":
        # The function counts a bin only if its magnitude rose against the lagged frame: a discrete decision on
        # float32 magnitudes.  The device's float64 FFT and scipy's agree to one float32 ulp per bin, so a bin whose
        # magnitude equals its lagged one to within that ulp goes either way (about one bin in 1e7; the reference's
        # own result moves the same way between its pyfftw and scipy backends).  3 of 80 runs of an 880-seed sweep
        # had one or two such rows: rows may differ by a bin or two, and such rows must be rare.
        bad = np.abs(got - want) > 1e-5 + 1e-4 * np.abs(want)
        assert bad.mean() <= 0.01, (what, int(bad.sum()))
        assert np.all(np.abs(got - want)[bad] <= 2 * np.abs(st).max(axis=1)[bad]), what
        return
    assert_close(got, want, atol=1e-5 * scale, what=what)
