"""Pin the numpy oracle (oracle/frontend.py) against
  (a) the reference's own known-answer literals (file:line cited per test), and
  (b) golden vectors produced by the real reference (tests/golden/make_golden.py).
CPU only."""
import numpy as np
import pytest

from oracle import frontend as O
from conftest import assert_close


# ------------------------------------------------------------------ framing (integers: bit-exact)
def test_num_frames_table(ref_framing):
    for n, hop, end, want in ref_framing["num_frames_cases"]:
        got = O.num_frames(int(n), hop, "extend" if end else "normal")
        assert got == int(want), (n, hop, end)


def test_num_frames_sample(sample_pcm, ref_sample):
    # reference: tests/test_audio_signal.py:1144-1164 (281 / 561 frames)
    assert O.num_frames(len(sample_pcm), 441.) == 281 == int(ref_sample["num_frames_100"])
    assert O.num_frames(len(sample_pcm), 44100 / 200.) == 561 == int(ref_sample["num_frames_200"])


def test_frame_cases(ref_framing):
    ramp = (np.arange(20000) % 30000 + 1).astype(np.int16)
    for N, hop, origin, idx, nf, first_nz, first_val, nnz, total in ref_framing["frame_cases"]:
        N, origin, idx = int(N), int(origin), int(idx)
        assert O.num_frames(len(ramp), hop) == int(nf)
        frame = O.signal_frame(ramp, idx, N, hop, origin)
        nzp = np.nonzero(frame)[0]
        assert len(nzp) == int(nnz)
        if len(nzp):
            assert nzp[0] == int(first_nz) and frame[nzp[0]] == int(first_val)
        assert int(frame.astype(np.int64).sum()) == int(total)


def test_origin_literals():
    # reference: tests/test_audio_signal.py:1265-1290 (origins 1023 / -1024 for frame_size 2048)
    assert O.resolve_origin("online", 2048) == 1023
    assert O.resolve_origin("stream", 2048) == -1024
    assert O.resolve_origin("offline", 2048) == 0


def test_frames_rows(sample_pcm, ref_sample):
    fr = O.framed(sample_pcm, 2048, 441.)
    assert fr.shape == (281, 2048)
    assert np.array_equal(fr[ref_sample["rows"]], ref_sample["frames_rows"])
    # reference: tests/test_audio_signal.py:1231-1232 -- frame 3 starts at sample 299
    assert O.frame_start(3, 2048, 441.) == 299


def test_diff_frames(ref_framing):
    for N, hop, ratio, want in ref_framing["diff_frames_cases"]:
        assert O.diff_frames(ratio, hop, int(N)) == int(want)


# ------------------------------------------------------------------ STFT / magnitude
def test_stft_small_known_answers():
    # reference: tests/test_audio_stft.py:57-101 (exact 12-point DFTs, circular shift sign flip, Nyquist)
    sig_2d = np.array([[1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0],
                       [1, 0, 0, 1, 0, 0, 1, 0, 0, 1, 0, 0],
                       [1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0]])
    res = O.stft(sig_2d, window=None)
    assert res.dtype == np.complex64
    assert np.allclose(res[0], [3, 0, 0, 3, 0, 0])
    assert np.allclose(res[1], [4, 0, 0, 0, 4, 0])
    assert np.allclose(res[2], [6, 0, 0, 0, 0, 0])
    res = O.stft(sig_2d, window=None, circular_shift=True)
    assert np.allclose(res[0], [3, 0, 0, -3, 0, 0])
    assert np.allclose(res[1], [4, 0, 0, 0, 4, 0])
    res = O.stft(sig_2d, window=None, include_nyquist=True)
    assert res.shape == (3, 7) and np.allclose(res[:, -1], [3, 0, 6])
    res = O.stft(sig_2d, window=None, fft_size=25)
    assert res.shape == (3, 12) and np.allclose(res[:, 0], [3, 4, 6])


def test_local_group_delay_known_answers():
    # reference: tests/test_audio_stft.py:141-147
    data = np.arange(20).reshape(10, 2) * 2
    assert np.allclose(np.tile([-2, 0], 10).reshape(10, 2), O.local_group_delay(data.astype(float)))
    data = np.arange(20).reshape(10, 2) * 4
    assert np.allclose(np.tile([2.28318531, 0], 10).reshape(10, 2), O.local_group_delay(data.astype(float)))


def test_stft_rows(sample_pcm, ref_sample):
    fr = O.framed(sample_pcm, 2048, 441.)
    fw = O.fft_window(np.hanning, 2048, sample_pcm.dtype)
    assert np.array_equal(fw, ref_sample["fft_window"])
    rows = ref_sample["rows"]
    X = O.stft(fr[rows], fw)
    assert np.array_equal(X, ref_sample["stft_rows"])
    Xc = O.stft(fr[rows], fw, circular_shift=True)
    assert np.array_equal(Xc, ref_sample["stft_circ_rows"])
    assert np.array_equal(O.spec(X), ref_sample["spec_rows"])


def test_spec_literals(sample_pcm):
    # reference: tests/test_audio_spectrogram.py:82-96
    fr = O.framed(sample_pcm, 2048, 441.)
    S = O.spec(O.stft(fr[:1], O.fft_window(np.hanning, 2048, np.int16)))
    assert np.allclose(S[0, :8], [3.15249, 4.00272, 5.66156, 6.30141, 6.02199, 10.84909, 17.83130, 19.44511])
    assert np.allclose(S[0, -8:], [0.0365325, 0.036513, 0.0364213, 0.0366203, 0.036737, 0.036423, 0.036335,
                                   0.0367054])


# ------------------------------------------------------------------ filterbank
def _compact(fb):
    starts, lens, w = [], [], []
    for b in range(fb.shape[1]):
        nz = np.nonzero(fb[:, b])[0]
        starts.append(nz[0]); lens.append(len(nz)); w.append(fb[nz, b])
    return np.array(starts), np.array(lens), np.concatenate(w)


@pytest.mark.parametrize("N", [1024, 2048, 4096])
@pytest.mark.parametrize("bpo,norm", [(6, True), (12, False), (12, True), (24, False)])
def test_filterbank_vs_reference(ref_filterbanks, N, bpo, norm):
    key = "N%d_b%d_%s" % (N, bpo, "norm" if norm else "raw")
    binf = O.fft_frequencies(N // 2, 44100)
    fb = O.logarithmic_filterbank(binf, bpo, 30., 17000., 440., norm, True)
    s, l, w = _compact(fb)
    assert np.array_equal(s, ref_filterbanks[key + "_start"])
    assert np.array_equal(l, ref_filterbanks[key + "_len"])
    assert np.array_equal(w, ref_filterbanks[key + "_w"])          # float32 weights bit-exact
    assert np.array_equal(O.filterbank_center_frequencies(fb, binf), ref_filterbanks[key + "_center"])


def test_filterbank_band_counts():
    # SURVEY.md section 8 table [probe]: 6/12/24 bands per octave
    want = {1024: (39, 69, 116), 2048: (45, 81, 140), 4096: (49, 91, 162)}
    for N, counts in want.items():
        binf = O.fft_frequencies(N // 2, 44100)
        got = tuple(O.logarithmic_filterbank(binf, b, norm_filters=(b == 6)).shape[1] for b in (6, 12, 24))
        assert got == counts


def test_filterbank_nonunique(ref_filterbanks):
    fb = O.logarithmic_filterbank(O.fft_frequencies(1024, 44100), 12, unique_filters=False)
    assert np.array_equal(fb, ref_filterbanks["N2048_b12_norm_nonunique_dense"])


def test_log_frequencies_literals():
    # reference: tests/test_audio_filters.py:224-233
    assert len(O.log_frequencies(6, 30, 17000)) == 55
    assert np.allclose(O.log_frequencies(1, 100, 1000), [110., 220., 440., 880.])
    assert np.allclose(O.log_frequencies(1, 100, 1000, 441), [110.25, 220.5, 441., 882.])


def test_frequencies2bins_literals():
    # reference: tests/test_audio_filters.py:396-421
    f2b = O.frequencies2bins
    assert np.allclose(f2b([0, 1, 1.1, 1.2, 2, 3, 3.5], [0, 1, 2, 3, 4]), [0, 1, 1, 1, 2, 3, 4])
    assert np.allclose(f2b([0, 1, 2, 3, 11], [0, 1, 2, 3, 4]), [0, 1, 2, 3, 4])
    assert np.allclose(f2b([-1, 0, 1, 2, 3, 4], [0, 1, 2, 3, 4]), [0, 0, 1, 2, 3, 4])
    assert np.allclose(f2b([0, 1, 2, 3, 4, 5], [0, 2, 4]), [0, 1, 1, 2, 2, 2])
    assert np.allclose(f2b([0, 1, 2, 3, 4, 5], [0, 2, 4], unique_bins=True), [0, 1, 2])


def test_triangular_filter_literals():
    # reference: tests/test_audio_filters.py:512-559
    tf = O.triangular_filter
    assert np.allclose(tf(0, 1, 2, False), [0, 1])
    assert np.allclose(tf(1, 2, 4, False), [0, 1, 0.5])
    assert np.allclose(tf(1, 2, 4, True), [0, 0.66667, 0.33333])
    assert np.allclose(tf(4, 6, 9, True), [0, 0.2, 0.4, 0.266667, 0.133333])
    assert np.allclose(tf(0, 0, 1, True), [1])
    assert list(O.triangular_band_bins(np.asarray([0, 1, 2, 3, 4, 6, 9]))) == \
        [(0, 1, 2), (1, 2, 3), (2, 3, 4), (3, 4, 6), (4, 6, 9)]
    with pytest.raises(ValueError):
        tf(0, 2, 1, False)


def test_logarithmic_filterbank_literals():
    # reference: tests/test_audio_filters.py:1051-1074
    fb = O.logarithmic_filterbank(np.arange(0, 20000, 20), 12)
    assert fb.shape == (1000, 81) and np.allclose(fb.min(), 0) and np.allclose(fb.max(), 1)
    fb = O.logarithmic_filterbank(np.arange(0, 20000, 20), 12, unique_filters=False)
    assert fb.shape == (1000, 108)
    binf = O.fft_frequencies(1024, 44100)
    fb = O.logarithmic_filterbank(binf)
    assert fb.shape == (1024, 81)
    nz = np.nonzero(fb)[0]
    assert np.allclose(binf[nz[0]], 43.066406) and np.allclose(binf[nz[-1]], 16731.298828)


# ------------------------------------------------------------------ filtered / log / diff on sample.wav
@pytest.fixture(scope="module")
def sample_stages(sample_pcm):
    return O.run_recipe(sample_pcm, O.SUPERFLUX, stages=True)


def test_filtered_log_default(sample_pcm, ref_sample):
    fr = O.framed(sample_pcm, 2048, 441.)
    S = O.spec(O.stft(fr, O.fft_window(np.hanning, 2048, np.int16)))
    fb = O.logarithmic_filterbank(O.fft_frequencies(1024, 44100))
    F = O.filtered(S, fb)
    assert F.shape == (281, 81)
    assert np.allclose(F, ref_sample["filt12n"], rtol=1e-6, atol=1e-7)
    # reference literals: tests/test_audio_spectrogram.py:146-154
    assert np.allclose(F[0, :8], [5.661564, 6.30141, 6.02199, 10.84909, 17.8313, 19.44511, 17.56456, 21.859523])
    L = O.logarithmic(F)
    assert np.allclose(L, ref_sample["logfilt12n"], rtol=1e-6, atol=1e-7)
    # reference literals: tests/test_audio_spectrogram.py:364-372
    assert np.allclose(L[0, :8], [0.8235762, 0.863407, 0.8464602, 1.073685, 1.27488, 1.3105896, 1.2686847,
                                  1.359067])
    assert np.allclose(L[0, -8:], [0.05042794, 0.0490095, 0.05608485, 0.05189138, 0.04567042, 0.03483925,
                                   0.02294769, 0.02719229])
    # reference literals: tests/test_audio_spectrogram.py:285-293 (log of the unfiltered spectrogram)
    LS = O.logarithmic(S)
    assert np.allclose(LS[0, :8], [0.618309, 0.699206, 0.823576, 0.86341, 0.84646, 1.073685, 1.27488, 1.310589])
    # reference literals: tests/test_audio_spectrogram.py:459-468 (signed diff of the plain spectrogram, d=1)
    D = O.difference(S, O.diff_frames(0.5, 441., 2048))
    assert np.allclose(D[1, :8], [1.13179708, -1.1511457, 2.7810955, 2.39441729, -4.87367058, -0.90269375,
                                  3.48209763, 11.14723015])
    assert np.allclose(D[1, -8:], [-0.01463442, -0.01408007, -0.01462659, -0.01431422, -0.01404046,
                                   -0.01457103, -0.01443923, -0.01443416])


def test_superflux_matrix(sample_stages, ref_sample):
    D = sample_stages["diff"]
    assert sample_stages["diff_frames"] == 1
    assert D.shape == (281, 140)
    assert np.allclose(sample_stages["log"], ref_sample["logfilt24"], rtol=1e-6, atol=1e-7)
    assert np.allclose(D, ref_sample["superflux_keepdims"], rtol=1e-6, atol=1e-6)
    assert np.allclose(D, ref_sample["superflux_diff"], rtol=1e-6, atol=1e-6)
    # reference literals: tests/test_audio_spectrogram.py:584-591
    assert np.allclose(D[1, :8], [0.11168772, 0.12317812, 0, 0, 0.03797626, 0.18899226, 0, 0.0903399])
    assert np.allclose(D[1, -8:], [0, 0, 0.01419619, 0, 0.02666602, 0.04325962, 0.10899737, 0.06546581])
    assert not D[0].any()
    assert np.allclose(sample_stages["odf"], ref_sample["odf_superflux_100"], rtol=1e-5, atol=1e-5)


def test_difference_processor_equals_keepdims(sample_stages, ref_sample):
    L = sample_stages["log"]
    P = O.difference_processor(L, 1, 3, True)
    assert np.array_equal(P, sample_stages["diff"])
    st = O.difference_processor(L, 1, 3, True, stack_diffs=np.hstack)
    assert st.shape == (281, 280)
    assert np.array_equal(st[:, :140], L) and np.array_equal(st[:, 140:], P)


def test_signed_diff_d2(sample_stages, ref_sample):
    D = O.difference(sample_stages["log"], 2, 3, False)
    assert np.allclose(D, ref_sample["diff_signed_d2"], rtol=1e-6, atol=1e-6)


# ------------------------------------------------------------------ ODFs, programs, peak picking
def test_logfiltspecflux(sample_pcm, ref_sample):
    st = O.run_recipe(sample_pcm, O.LOGFILTSPECFLUX, stages=True)
    assert np.allclose(st["log"], ref_sample["logfilt12"], rtol=1e-6, atol=1e-7)
    assert np.allclose(st["odf"], ref_sample["odf_logfiltspecflux"], rtol=1e-5, atol=1e-5)
    # the reference's shipped fixture (tests/test_bin.py:702-738, atol 1e-5)
    assert np.allclose(st["odf"], ref_sample["fixture_act_log_filt_spec_flux"], atol=1e-5)
    on = O.onset_peak_picking(st["odf"], 100, threshold=1.6, pre_max=0.01, post_max=0.05, pre_avg=0.15,
                              combine=0.03)
    assert np.array_equal(on, ref_sample["onsets_logfiltspecflux"])
    assert np.allclose(on, ref_sample["fixture_det_log_filt_spec_flux"], atol=1e-5)


def test_superflux_program_200(sample_pcm, ref_sample):
    st = O.run_recipe(sample_pcm, O.SUPERFLUX_200, stages=True)
    assert st["diff_frames"] == 2 and st["odf"].shape == (561,)
    assert np.allclose(st["odf"], ref_sample["odf_superflux_200"], rtol=1e-5, atol=1e-5)
    assert np.allclose(st["odf"], ref_sample["fixture_act_super_flux"], atol=1e-5)   # tests/test_bin.py:956-1003
    peaks = O.peak_picking(st["odf"], 1.1, None, 30, 0, 2, 10)
    assert np.array_equal(peaks, ref_sample["peaks_superflux_200"])
    on = O.onset_peak_picking(st["odf"], 200, threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15,
                              combine=0.03)
    assert np.array_equal(on, ref_sample["onsets_superflux_200"])
    assert np.allclose(on, ref_sample["fixture_det_super_flux"], atol=1e-5)


def test_complex_flux_200(sample_pcm, ref_sample):
    r = dict(O.SUPERFLUX_200, circular_shift=True)
    st = O.run_recipe(sample_pcm, r, stages=True)
    lgd = O.local_group_delay(O.phase(st["stft"]))
    assert np.allclose(lgd[ref_sample["rows"]], ref_sample["lgd_200_rows"], atol=1e-5)
    odf = O.complex_flux(st["log"], st["stft"], st["filterbank"], st["diff_frames"])
    assert np.allclose(odf, ref_sample["odf_complexflux_200"], rtol=1e-5, atol=1e-5)
    assert np.allclose(odf, ref_sample["fixture_act_complex_flux"], atol=1e-5)        # tests/test_bin.py:350-386
    on = O.onset_peak_picking(odf, 200, threshold=0.25, pre_max=0.01, post_max=0.05, pre_avg=0.15,
                              combine=0.03)
    assert np.array_equal(on, ref_sample["onsets_complexflux_200"])
    assert np.allclose(on, ref_sample["fixture_det_complex_flux"], atol=1e-5)


def test_peak_picking_literals(ref_sample):
    # reference: tests/test_features_onsets.py:233-245 on the shipped fixture sample.super_flux.npz (fps 200)
    act = ref_sample["fixture_act_super_flux"]
    onsets = O.peak_picking(act, 1.1)
    assert list(onsets[:6]) == [2, 10, 17, 48, 55, 80] and len(onsets) == 35
    assert list(O.peak_picking(act, 1.1, smooth=3)[:6]) == [2, 10, 17, 24, 48, 55]
    assert list(O.peak_picking(act, 1.1, pre_max=2, post_max=10, pre_avg=30)[:6]) == [2, 17, 55, 89, 122, 159]
    # reference: tests/test_features_onsets.py:262-287 (OnsetPeakPickingProcessor, 16 onset times)
    on = O.onset_peak_picking(act, 200, threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0,
                              combine=0.03, delay=0)
    assert np.allclose(on, [0.01, 0.085, 0.275, 0.445, 0.61, 0.795, 0.98, 1.115, 1.365, 1.475, 1.62, 1.795,
                            2.14, 2.33, 2.485, 2.665])


def test_odf_literals(sample_pcm):
    # reference: tests/test_features_onsets.py:24-27,62-86: LogarithmicFilteredSpectrogram(Spectrogram(
    # sample, circular_shift=True), num_bands=24, mul=1, add=1) -> norm_filters defaults to True
    r = dict(frame_size=2048, hop_size=441., num_bands=24, norm_filters=True, circular_shift=True,
             diff_max_bins=None, positive_diffs=True)
    st = O.run_recipe(sample_pcm, r, stages=True)
    assert np.allclose(st["odf"][:6], [0, 3.91207361, 2.91675663, 1.38361311, 2.59582925, 2.16986609])
    sf = O.flux(O.difference(st["log"], st["diff_frames"], 3, True))
    assert np.allclose(sf[:6], [0, 2.08680153, 0.6411702, 0.38634294, 0.40202433, 0.63349575])
    cf = O.complex_flux(st["log"], st["stft"], st["filterbank"], st["diff_frames"])
    assert np.allclose(cf[:6], [0, 0.476213485, 0.0877621323, 0.0593151376, 0.0654867291, 0.0954693183])


def _odf_inputs(x):
    r = dict(frame_size=2048, hop_size=441., num_bands=24, norm_filters=True, circular_shift=True,
             diff_max_bins=None, positive_diffs=True)
    return O.run_recipe(x, r, stages=True)


def _oracle_odfs(st):
    d = st["diff_frames"]
    sp = O.spec(st["stft"])
    return {"hfc_logfilt": O.high_frequency_content(st["log"]), "hfc_spec": O.high_frequency_content(sp),
            "sd_logfilt": O.spectral_diff(st["log"], d), "sd_spec": O.spectral_diff(sp, d),
            "mkl_logfilt": O.modified_kullback_leibler(st["log"]), "mkl_spec": O.modified_kullback_leibler(sp),
            "pd": O.phase_deviation(st["stft"]), "wpd": O.weighted_phase_deviation(st["stft"]),
            "nwpd": O.normalized_weighted_phase_deviation(st["stft"]), "cd": O.complex_domain(st["stft"]),
            "rcd": O.rectified_complex_domain(st["stft"], d)}


def test_other_odfs_literals(sample_pcm):
    # reference known answers: tests/test_features_onsets.py:35-152 (np.allclose defaults)
    o = _oracle_odfs(_odf_inputs(sample_pcm))
    want = {"hfc_logfilt": [8.97001563, 9.36399107, 8.64144536, 8.34977449, 8.21097918, 8.40412515],
            "sd_logfilt": [0, 0.55715936, 0.64004618, 0.0810971, 0.295396, 0.16324584],
            "mkl_logfilt": [0, 0.71910584, 0.6664055, 0.68092251, 0.69984031, 0.71744561],
            "pd": [0, 0, 0.71957183, 0.91994524, 0.9418999, 0.86083585],
            "wpd": [0, 0, 0.19568817, 0.20483065, 0.17890805, 0.16970603],
            "nwpd": [0, 0, 0.46018526, 0.50193471, 0.42031503, 0.40806249],
            "cd": [399.29980469, 585.9564209, 262.08010864, 225.84718323, 196.88954163, 200.32469177],
            "rcd": [0, 394.165222, 119.79425, 96.70564, 122.52311, 92.61698]}
    for k, v in want.items():
        # the reference literals were produced with a pyfftw float32-planned FFT upstream; the scipy float64
        # path reproduces them to 1e-5 relative (same margin as tests/test_bin.py uses for the fixtures)
        assert np.allclose(o[k][:6], v, rtol=2e-5, atol=1e-6), (k, o[k][:6])


@pytest.mark.parametrize("which", ["sample", "synth"])
def test_other_odfs_golden(sample_pcm, ref_odfs, which):
    from madmom_b200.synth import make_track
    x = sample_pcm if which == "sample" else make_track(3, 2.0)
    o = _oracle_odfs(_odf_inputs(x))
    for k, v in o.items():
        want = ref_odfs["%s_%s" % (which, k)]
        assert v.dtype == want.dtype and v.shape == want.shape, k
        assert np.allclose(v, want, rtol=1e-6, atol=1e-7), (k, np.abs(v - want).max())


# ------------------------------------------------------------------ multi-resolution stack, synthetic tracks
def multires_oracle(x, mul, diff_ratio, hop=441.):
    cols = []
    for N in (1024, 2048, 4096):
        st = O.run_recipe(x, dict(frame_size=N, hop_size=hop, num_bands=6, norm_filters=True, mul=mul, add=1,
                                  diff_ratio=diff_ratio, positive_diffs=True), stages=True)
        cols += [st["log"], st["diff"]]
    return np.hstack(cols)


def test_multires_sample(sample_pcm, ref_sample):
    on = multires_oracle(sample_pcm, 5, 0.25)
    assert on.shape == (281, 266)
    assert np.allclose(on, ref_sample["rnn_onset_stack"], rtol=1e-6, atol=1e-6)
    be = multires_oracle(sample_pcm, 1, 0.5)
    assert np.allclose(be, ref_sample["rnn_beat_stack"], rtol=1e-6, atol=1e-6)


def test_synthetic_tracks(ref_synth):
    from madmom_b200.synth import make_track
    for seed, seconds in ((1000, 3.0), (1001, 2.5)):
        x = make_track(seed, seconds)
        tag = "s%d" % seed
        assert np.allclose(O.run_recipe(x, O.SUPERFLUX), ref_synth[tag + "_superflux"], rtol=1e-6, atol=1e-6)
        st = O.run_recipe(x, O.LOGFILTSPECFLUX, stages=True)
        assert np.allclose(st["log"], ref_synth[tag + "_lfsf_log"], rtol=1e-6, atol=1e-6)
        assert np.allclose(st["diff"], ref_synth[tag + "_lfsf_diff"], rtol=1e-6, atol=1e-6)
        assert np.allclose(st["odf"], ref_synth[tag + "_lfsf_odf"], rtol=1e-5, atol=1e-5)
        st = O.run_recipe(x, O.SUPERFLUX_200, stages=True)
        assert np.allclose(st["odf"], ref_synth[tag + "_superflux200_odf"], rtol=1e-5, atol=1e-5)
        assert np.allclose(multires_oracle(x, 5, 0.25), ref_synth[tag + "_rnn_onset"], rtol=1e-6, atol=1e-6)
        assert np.allclose(multires_oracle(x, 1, 0.5), ref_synth[tag + "_rnn_beat"], rtol=1e-6, atol=1e-6)
    xf = make_track(1000, 3.0, dtype=np.float32)
    assert np.allclose(O.run_recipe(xf, O.SUPERFLUX), ref_synth["s1000_f32_superflux"], rtol=1e-6, atol=1e-6)


def test_phase_and_local_group_delay_vs_reference(ref_phase):
    """oracle.phase / oracle.local_group_delay (madmom/audio/stft.py:136-179) on the reference's own STFT rows equal
    the reference's Phase / LocalGroupDelay rows; literals of tests/test_audio_stft.py:141-147."""
    from oracle import frontend as O
    ph = O.phase(ref_phase["stft_rows"])
    assert ph.dtype == np.float32 and np.array_equal(ph, ref_phase["phase_rows"])
    assert np.array_equal(O.local_group_delay(ph), ref_phase["lgd_rows"])
    assert np.array_equal(ref_phase["lgd_rows"], ref_phase["lgd_from_class_rows"])
    assert np.allclose(O.local_group_delay(np.arange(20).reshape(10, 2) * 2), np.tile([-2, 0], 10).reshape(10, 2))
    assert np.allclose(O.local_group_delay(np.arange(20).reshape(10, 2) * 4),
                       np.tile([2.28318531, 0], 10).reshape(10, 2))


def test_peak_picking_rnn_fixture(ref_acts):
    """The oracle's peak picking against the real reference on its sample.onsets_rnn.npz fixture
    (tests/golden/make_golden_acts.py) and against the literals of tests/test_features_onsets.py:247-269."""
    act, fps = ref_acts["rnn_act"], float(ref_acts["rnn_fps"])
    peaks = O.peak_picking(act, 0.23, post_max=0)
    assert np.array_equal(peaks, ref_acts["rnn_peaks_online_params"])
    assert list(peaks) == [1, 3, 10, 12, 29, 46, 62, 63, 77, 79, 81, 99, 100, 113, 115, 148, 149, 164, 181, 183, 216,
                           234, 250, 268]
    sf, sf_fps = ref_acts["sf_act"], float(ref_acts["sf_fps"])
    kw = dict(threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15, post_avg=0, combine=0.03)
    got = O.onset_peak_picking(sf, sf_fps, delay=0, **kw)
    assert np.array_equal(got, ref_acts["sf_onsets"])
    assert np.allclose(got, [0.01, 0.085, 0.275, 0.445, 0.61, 0.795, 0.98, 1.115, 1.365, 1.475, 1.62, 1.795, 2.14,
                             2.33, 2.485, 2.665])
    assert np.array_equal(O.onset_peak_picking(sf, sf_fps, delay=1, **kw), ref_acts["sf_onsets_delay1"])
    # the online processor fed everything at once equals the offline one with the parameters it ends up with
    off = O.onset_peak_picking(act, fps, threshold=0.23, smooth=0, post_avg=0, post_max=0)
    assert np.array_equal(off, ref_acts["rnn_onsets_offline_same_params"])
    assert np.array_equal(off, ref_acts["rnn_onsets_online"])
    assert np.allclose(off, [0.01, 0.1, 0.29, 0.46, 0.62, 0.77, 0.81, 0.99, 1.13, 1.48, 1.64, 1.81, 2.16, 2.34, 2.5,
                             2.68])
    for combine in (0.0, 0.03, 0.1):
        got = O.onset_peak_picking(act, fps, threshold=0.02, smooth=0.02, pre_avg=0.1, post_avg=0.05, pre_max=0.02,
                                   post_max=0.03, combine=combine, delay=0.005)
        assert np.array_equal(got, ref_acts["rnn_onsets_combine_%g" % combine]), combine
