"""GPU parity tests of the C-ABI path (libmadmom_b200.so through madmom_b200._lib.FrontEndPlan) against the
CPU oracle on the same inputs.  Bounds: integer outputs bit-exact; float32 features
|gpu - ref| <= 1e-5 + 1e-4 |ref|."""
import numpy as np
import pytest

from conftest import assert_close

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _plan_for(x_dtype, N, hop, bpo, norm, mul=1.0, add=1.0, ratio=0.5, max_bins=None, positive=True, circ=False,
              origin=0, chunk_frames=0):
    from oracle import frontend as O
    from madmom_b200 import _lib
    win = np.hanning(N)
    fw = O.fft_window(win, N, x_dtype)
    fb = O.logarithmic_filterbank(O.fft_frequencies(N // 2, 44100), bpo, 30., 17000., 440., norm, True)
    d = O.diff_frames(ratio, hop, N, win)
    plan = _lib.FrontEndPlan(N, hop, fw, pcm_dtype=x_dtype, origin=origin, circular_shift=circ, filterbank=fb,
                             log_mode=_lib.LOG_LOG10, mul=mul, add=add, diff_frames=d, diff_max_bins=max_bins or 0,
                             positive_diffs=positive, chunk_frames=chunk_frames)
    recipe = dict(frame_size=N, hop_size=hop, origin=origin, num_bands=bpo, norm_filters=norm, mul=mul, add=add,
                  diff_ratio=ratio, diff_max_bins=max_bins, positive_diffs=positive, circular_shift=circ)
    return plan, recipe


def _run_all(plan, tracks, hop):
    """tracks: list of 1-D numpy arrays.  Returns dict of numpy outputs for the concatenated batch."""
    from madmom_b200 import _lib
    lens = [len(t) for t in tracks]
    T = [_lib.num_frames(n, hop) for n in lens]
    units = _lib.make_units(lens, T)
    pcm = torch.from_numpy(np.concatenate(tracks)).cuda()
    rows, K, B = int(sum(T)), plan.num_bins, plan.num_bands
    outs = {"stft": torch.full((rows, K), float("nan"), dtype=torch.complex64, device="cuda"),
            "spec": torch.full((rows, K), float("nan"), device="cuda"),
            "filt": torch.full((rows, B), float("nan"), device="cuda"),
            "log": torch.full((rows, B), float("nan"), device="cuda"),
            "diff": torch.full((rows, B), float("nan"), device="cuda"),
            "odf": torch.full((rows,), float("nan"), device="cuda")}
    plan.run(pcm, units, outs)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in outs.items()}, T


def assert_bins_close(got, ref, what=""):
    """Raw bins / magnitudes: the reference rounds a float64 FFT once to complex64 (stft.py:26, :105-131) and so does
    the float64 kernel (csrc/stft64.cu): per component at most one float32 ulp (rounding of two float64 results that
    differ in their last bits) plus 1e-10 of the largest bin."""
    got, ref = np.asarray(got), np.asarray(ref)
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    scale = float(np.abs(ref).max()) if ref.size else 0.0
    parts = [(got.real, ref.real), (got.imag, ref.imag)] if np.iscomplexobj(ref) else [(got, ref)]
    for a, b in parts:
        err = np.abs(a.astype(np.float64) - b.astype(np.float64))
        bound = 3e-7 * np.abs(b.astype(np.float64)) + 1e-10 * scale
        assert (err <= bound).all(), "%s: worst excess %g (scale %g)" % (what, float((err - bound).max()), scale)


def _check_stages(got, ref, what):
    assert_bins_close(got["stft"], ref["stft"], what + " stft")
    assert_bins_close(got["spec"], ref["spec"], what + " spec")
    assert_close(got["filt"], ref["filt"], what=what + " filt")
    assert_close(got["log"], ref["log"], what=what + " log")
    assert_close(got["diff"], ref["diff"], what=what + " diff")
    assert_close(got["odf"], ref["odf"], what=what + " odf")


@pytest.mark.parametrize("N", [2048, 1024, 4096])
def test_sample_wav_all_stages(sample_pcm, N):
    from oracle import frontend as O
    plan, recipe = _plan_for(np.int16, N, 441., 24, False, max_bins=3)
    got, T = _run_all(plan, [sample_pcm], 441.)
    assert T == [281]
    ref = O.run_recipe(sample_pcm, recipe, stages=True)
    _check_stages(got, ref, "sample.wav N=%d" % N)
    assert not got["diff"][:ref["diff_frames"]].any()


def test_superflux_golden(sample_pcm, ref_sample):
    """Config 1 against the vectors produced by the real reference (tests/golden/make_golden.py) and the
    reference's own literals (tests/test_audio_spectrogram.py:584-591)."""
    plan, _ = _plan_for(np.int16, 2048, 441., 24, False, max_bins=3)
    got, _ = _run_all(plan, [sample_pcm], 441.)
    assert_close(got["diff"], ref_sample["superflux_diff"], what="superflux vs reference")
    assert_close(got["log"], ref_sample["logfilt24"], what="logfilt24 vs reference")
    assert_close(got["odf"], ref_sample["odf_superflux_100"], what="odf vs reference")
    assert np.allclose(got["diff"][1, :8], [0.11168772, 0.12317812, 0, 0, 0.03797626, 0.18899226, 0, 0.0903399],
                       atol=1e-5)
    assert np.allclose(got["spec"][0, :8], [3.15249, 4.00272, 5.66156, 6.30141, 6.02199, 10.84909, 17.83130,
                                            19.44511])
    rows = ref_sample["rows"]
    assert_bins_close(got["stft"][rows], ref_sample["stft_rows"], "stft rows vs reference")
    assert_bins_close(got["spec"][rows], ref_sample["spec_rows"], "spec rows vs reference")


def test_fps200_fractional_hop_and_peaks(sample_pcm, ref_sample):
    """bin/SuperFlux recipe: hop 220.5 (frame starts int(i*220.5) must be exact), d=2; onset indices from the
    GPU ODF must equal the reference's, bit-exact."""
    from oracle import frontend as O
    hop = 44100 / 200.
    plan, recipe = _plan_for(np.int16, 2048, hop, 24, False, max_bins=3)
    got, T = _run_all(plan, [sample_pcm], hop)
    assert T == [561] and plan.diff_frames == 2
    assert_close(got["odf"], ref_sample["odf_superflux_200"], what="superflux fps200")
    assert np.allclose(got["odf"], ref_sample["fixture_act_super_flux"], atol=1e-5)   # tests/test_bin.py:956-1003
    peaks = O.peak_picking(got["odf"], 1.1, None, 30, 0, 2, 10)
    assert np.array_equal(peaks, ref_sample["peaks_superflux_200"])
    on = O.onset_peak_picking(got["odf"], 200, threshold=1.1, pre_max=0.01, post_max=0.05, pre_avg=0.15,
                              combine=0.03)
    assert np.array_equal(on, ref_sample["onsets_superflux_200"])


def test_circular_shift_stft(sample_pcm, ref_sample):
    plan, _ = _plan_for(np.int16, 2048, 441., 24, False, circ=True)
    got, _ = _run_all(plan, [sample_pcm], 441.)
    rows = ref_sample["rows"]
    assert_bins_close(got["stft"][rows], ref_sample["stft_circ_rows"], "circular shift vs reference")


@pytest.mark.parametrize("N", [1024, 2048, 4096])
def test_include_nyquist(sample_pcm, N):
    """include_nyquist (stft.py:93-95): frame_size / 2 + 1 bins for the stft / spec outputs, float32 PCM and a
    circular shift included; a plan with a filterbank refuses it."""
    from oracle import frontend as O
    from madmom_b200 import _lib
    x = (sample_pcm[:30000].astype(np.float64) / 32767.).astype(np.float32) if N == 1024 else sample_pcm[:30000]
    circ = N == 4096
    fw = O.fft_window(np.hanning(N), N, x.dtype)
    plan = _lib.FrontEndPlan(N, 441., fw, pcm_dtype=x.dtype, circular_shift=circ, include_nyquist=True)
    assert plan.num_bins == N // 2 + 1
    T = _lib.num_frames(len(x), 441.)
    outs = {"stft": torch.empty((T, N // 2 + 1), dtype=torch.complex64, device="cuda"),
            "spec": torch.empty((T, N // 2 + 1), device="cuda")}
    assert plan.run(torch.from_numpy(x).cuda(), _lib.make_units([len(x)], [T]), outs) == 1
    ref = O.stft(O.framed(x, N, 441.), fw, circular_shift=circ, include_nyquist=True)
    assert ref.shape == (T, N // 2 + 1)
    assert_bins_close(outs["stft"].cpu().numpy(), ref, "include_nyquist stft N=%d" % N)
    assert_bins_close(outs["spec"].cpu().numpy(), O.spec(ref), "include_nyquist spec N=%d" % N)
    with pytest.raises(NotImplementedError):
        _lib.FrontEndPlan(N, 441., fw, pcm_dtype=x.dtype, include_nyquist=True,
                          filterbank=O.logarithmic_filterbank(O.fft_frequencies(N // 2, 44100), 12))


@pytest.mark.parametrize("N,bpo,norm,mul,ratio", [(2048, 12, False, 1., .5), (1024, 6, True, 5., .25),
                                                   (4096, 6, True, 5., .25), (4096, 6, True, 1., .5)])
def test_ragged_batch_vs_oracle(N, bpo, norm, mul, ratio):
    """Batch of ragged synthetic tracks (incl. shorter than one frame / one hop) in one launch; small chunks
    so that chunk boundaries, lead-in recomputation and track edges are all exercised."""
    from oracle import frontend as O
    from madmom_b200.synth import make_track
    lens = [44100, 1, 440, 441, 442, 30011, 2048, 5000, 77777]
    tracks = [make_track(1000 + i, num_samples=n) for i, n in enumerate(lens)]
    plan, recipe = _plan_for(np.int16, N, 441., bpo, norm, mul=mul, ratio=ratio, chunk_frames=13)
    got, T = _run_all(plan, tracks, 441.)
    r0 = 0
    for x, t in zip(tracks, T):
        ref = O.run_recipe(x, recipe, stages=True)
        sl = {k: v[r0:r0 + t] for k, v in got.items()}
        _check_stages(sl, ref, "track len %d N=%d" % (len(x), N))
        r0 += t


@pytest.mark.parametrize("N,ratio", [(1024, 0.25), (1024, 0.5), (2048, 0.5), (4096, 0.25)])
def test_rows_do_not_depend_on_chunking(N, ratio):
    """A track's rows are bit-identical whatever the chunk size (the library's default, its choice for a small
    launch, explicit odd / even sizes) and whether the track runs alone or inside a batch.  N = 1024 packs two
    frames into one FFT: pairs are formed on the global frame index, not on the position inside a chunk."""
    from madmom_b200.synth import make_track
    lens = [50000, 44100 * 4 + 17, 441 * 37, 12345]
    tracks = [make_track(2000 + i, num_samples=n) for i, n in enumerate(lens)]
    ref = None
    for chunk in (0, 13, 14, 50, 333):
        plan, _ = _plan_for(np.int16, N, 441., 6, True, ratio=ratio, chunk_frames=chunk)
        got, T = _run_all(plan, tracks, 441.)
        if ref is None:
            ref = got
            alone, _ = _run_all(plan, tracks[1:2], 441.)       # a small launch: fewer rounds per chunk
            for k in ("log", "diff", "odf", "spec"):
                assert np.array_equal(alone[k], got[k][T[0]:T[0] + T[1]], equal_nan=True), (N, k)
        for k in ("log", "diff", "odf", "spec", "stft"):
            assert np.array_equal(got[k], ref[k], equal_nan=True), (N, chunk, k)


@pytest.mark.parametrize("N", [1024, 2048, 4096])
def test_rows_do_not_depend_on_the_work_list(N, monkeypatch):
    """The default cut (equal rounds per warp group, in waves: work_list.h) against fixed chunks, for a batch large
    enough that every CTA gets several waves, with wave lengths that put the cuts everywhere."""
    from madmom_b200.synth import make_track
    rng = np.random.default_rng(N)
    lens = [int(v) for v in rng.integers(1, 44100 * 12, 40)] + [0, 1, 441]
    tracks = [make_track(3000 + i, num_samples=n) for i, n in enumerate(lens)]
    plan, _ = _plan_for(np.int16, N, 441., 6, True, ratio=0.25, chunk_frames=50)
    ref, _ = _run_all(plan, tracks, 441.)
    for wave in ("0", "1", "3", "64"):
        monkeypatch.setenv("MMB_WAVE_ROUNDS", wave)
        plan, _ = _plan_for(np.int16, N, 441., 6, True, ratio=0.25)
        got, _ = _run_all(plan, tracks, 441.)
        for k in ("log", "diff", "odf"):
            assert np.array_equal(got[k], ref[k], equal_nan=True), (N, wave, k)


def _window_case(N, window, x, monkeypatch=None, table=False):
    """log / diff / odf of `x` for frame size N and the window array, from the library and from the oracle."""
    from oracle import frontend as O
    from madmom_b200 import _lib
    if monkeypatch is not None:
        if table:
            monkeypatch.setenv("MMB_WIN_TABLE", "1")
        else:
            monkeypatch.delenv("MMB_WIN_TABLE", raising=False)
    fw = O.fft_window(window, N, x.dtype)
    fb = O.logarithmic_filterbank(O.fft_frequencies(N // 2, 44100), 12, 30., 17000., 440., False, True)
    plan = _lib.FrontEndPlan(N, 441., fw, pcm_dtype=x.dtype, filterbank=fb, log_mode=_lib.LOG_LOG10, mul=1., add=1.,
                             diff_frames=1, diff_max_bins=0, positive_diffs=True)
    got, _ = _run_all(plan, [x], 441.)
    st = O.stft(O.framed(x, N, 441.), fw)
    log = O.logarithmic(O.filtered(O.spec(st), fb), 1., 1.)
    diff = O.difference(log, 1, None, True)
    return got, {"log": log, "diff": diff, "odf": O.flux(diff)}


@pytest.mark.parametrize("N", [2048, 4096])
@pytest.mark.parametrize("dtype", [np.int16, np.float32])
def test_cosine_sum_windows_are_evaluated_in_the_kernel(N, dtype, monkeypatch):
    """np.hanning / np.hamming (any scale) are recognised at plan creation and evaluated in the kernel instead of
    loaded (frame_fft.cuh: *_wcos): against the oracle within the feature bound, and against the table path
    (MMB_WIN_TABLE=1) far inside it (4e-6 per band) -- a window value moves by at most 3e-7 of the window's peak."""
    from madmom_b200.synth import make_track
    x = make_track(77, 2.0, dtype=dtype)
    for window in (np.hanning(N), np.hamming(N), 3.5 * np.hanning(N), 0.25 - 0.75 * np.cos(2 * np.pi * np.arange(N) / (N - 1))):
        got, ref = _window_case(N, window, x, monkeypatch, table=False)
        tab, _ = _window_case(N, window, x, monkeypatch, table=True)
        for k in ("log", "diff", "odf"):
            assert_close(got[k], ref[k], what="cosine-sum window N=%d %s" % (N, k))
            assert_close(tab[k], ref[k], what="table window N=%d %s" % (N, k))
            err = np.abs(got[k].astype(np.float64) - tab[k].astype(np.float64))
            bands = tab["log"].shape[1]
            bound = 4e-6 * (bands if k == "odf" else 1) + 4e-6 * np.abs(tab[k].astype(np.float64))
            assert (err <= bound).all(), (N, k, float(err.max()))
        assert not np.array_equal(got["log"], tab["log"])          # two paths indeed


@pytest.mark.parametrize("N", [2048, 4096])
def test_other_windows_keep_the_table(N, monkeypatch):
    """Windows that are not a cosine sum of the symmetric kind (Blackman, the periodic Hann, a rectangle, a Hann
    with one tap off by 1e-6) go through the table: same bits with and without MMB_WIN_TABLE, oracle bound met."""
    from madmom_b200.synth import make_track
    x = make_track(78, 1.5)
    n = np.arange(N)
    bumped = np.hanning(N)
    bumped[N // 3] += 1e-6
    for window in (np.blackman(N), 0.5 - 0.5 * np.cos(2 * np.pi * n / N), np.ones(N), bumped):
        got, ref = _window_case(N, window, x, monkeypatch, table=False)
        tab, _ = _window_case(N, window, x, monkeypatch, table=True)
        for k in ("log", "diff", "odf"):
            assert_close(got[k], ref[k], what="window N=%d %s" % (N, k))
            assert np.array_equal(got[k], tab[k]), (N, k)


def test_float32_pcm(ref_synth):
    from oracle import frontend as O
    from madmom_b200.synth import make_track
    x = make_track(1000, 3.0, dtype=np.float32)
    plan, recipe = _plan_for(np.float32, 2048, 441., 24, False, max_bins=3)
    got, _ = _run_all(plan, [x], 441.)
    assert_close(got["diff"], ref_synth["s1000_f32_superflux"], what="float32 pcm superflux")


def test_signed_diff_and_origin():
    from oracle import frontend as O
    from madmom_b200.synth import make_track
    x = make_track(5, 2.0)
    for origin in ("online", "stream", 100):
        o = O.resolve_origin(origin, 2048)
        plan, recipe = _plan_for(np.int16, 2048, 441., 12, True, positive=False, max_bins=3, origin=o)
        got, _ = _run_all(plan, [x], 441.)
        ref = O.run_recipe(x, recipe, stages=True)
        _check_stages(got, ref, "origin %s" % origin)


def test_frame_range_units_match_whole_signal():
    """Shard mode: frame ranges with a PCM halo reproduce the unsharded rows bit for bit."""
    from madmom_b200 import _lib
    from madmom_b200.synth import make_track
    x = make_track(9, 6.0)
    hop, N = 441., 2048
    plan, _ = _plan_for(np.int16, N, hop, 24, False, max_bins=3)
    whole, T = _run_all(plan, [x], hop)
    T = T[0]
    d = plan.diff_frames
    cuts = [0, 100, 333, T]
    units = np.zeros(3, dtype=_lib.TRACK_DTYPE)
    pieces, off = [], 0
    for i in range(3):
        f0, f1 = cuts[i], cuts[i + 1]
        s0 = max(_lib.frame_start(max(f0 - d, 0), N, hop, 0), 0)
        s1 = min(_lib.frame_start(f1 - 1, N, hop, 0) + N, len(x))
        pieces.append(x[s0:s1])
        units[i] = (off, s1 - s0, s0, f0, f1 - f0, f0)
        off += s1 - s0
    pcm = torch.from_numpy(np.concatenate(pieces)).cuda()
    outs = {"diff": torch.zeros((T, plan.num_bands), device="cuda"), "odf": torch.zeros((T,), device="cuda")}
    plan.run(pcm, units, outs)
    torch.cuda.synchronize()
    assert np.array_equal(outs["diff"].cpu().numpy(), whole["diff"])
    assert np.array_equal(outs["odf"].cpu().numpy(), whole["odf"])


def test_host_path_equals_device_path():
    from madmom_b200 import _lib
    from madmom_b200.synth import make_track
    tracks = [make_track(50 + i, num_samples=n) for i, n in enumerate([40000, 12345, 88200])]
    plan, _ = _plan_for(np.int16, 2048, 441., 12, False)
    dev, T = _run_all(plan, tracks, 441.)
    units = _lib.make_units([len(t) for t in tracks], T)
    rows = sum(T)
    pcm = np.concatenate(tracks)
    stacked = np.zeros((rows, 2 * plan.num_bands), np.float32)       # hstack((log, diff)) written in place
    odf = np.zeros(rows, np.float32)
    n = plan.run_host(pcm, units, {"log": stacked[:, :plan.num_bands], "diff": stacked[:, plan.num_bands:],
                                   "odf": odf})
    assert n >= 1
    assert np.array_equal(stacked[:, :plan.num_bands], dev["log"])
    assert np.array_equal(stacked[:, plan.num_bands:], dev["diff"])
    assert np.array_equal(odf, dev["odf"])


@pytest.mark.parametrize("N,ratio", [(2048, 0.5), (1024, 0.5), (4096, 0.25)])
def test_host_path_splits_a_long_unit(N, ratio, monkeypatch):
    """A unit larger than a pipeline group (one long file) is cut into frame ranges with their PCM halo inside
    mmb_run_host so that copies and kernels overlap; rows stay bit-identical to the device path (frame starts by the
    global index, lead-in frames recomputed; frame pairing of N = 1024 included).  A short unit before and after
    the long one, an odd number of samples, the group size forced down to 1 MiB."""
    from madmom_b200 import _lib
    from madmom_b200.synth import make_track
    tracks = [make_track(70 + i, num_samples=n) for i, n in enumerate([30001, 44100 * 100 + 77, 5000])]
    plan, _ = _plan_for(np.int16, N, 441., 6, True, ratio=ratio)
    dev, T = _run_all(plan, tracks, 441.)
    units = _lib.make_units([len(t) for t in tracks], T)
    rows = sum(T)
    pcm = np.concatenate(tracks)
    monkeypatch.setenv("MMB_HOST_GROUP_MB", "1")
    out = {"log": np.zeros((rows, plan.num_bands), np.float32), "diff": np.full((rows, plan.num_bands), np.nan, np.float32),
           "odf": np.zeros(rows, np.float32), "spec": np.zeros((rows, plan.num_bins), np.float32)}
    n = plan.run_host(pcm, units, out)
    assert n >= 8                                        # the 8.8 MB unit went through as several groups
    for k in out:
        assert np.array_equal(out[k], dev[k]), (N, k)


@pytest.mark.parametrize("seed", range(int(__import__("os").environ.get("MMB_FUZZ_SEEDS", "16"))))
def test_random_host_path_equals_device_path(seed, monkeypatch):
    """mmb_run_host (pinned staging, groups of units on three streams, over-long units cut into frame ranges) against
    the device path on random batches: ragged track lengths from one sample to a few megabytes, random group size,
    frame size, hop, lag, int16 / float32 -- every output bit-identical."""
    from madmom_b200 import _lib
    from madmom_b200.synth import make_track
    rng = np.random.default_rng(9000 + seed)
    N = int(rng.choice([1024, 2048, 4096]))
    hop = float(rng.choice([441., 220.5, 882., 300.25]))
    dtype = np.float32 if rng.integers(3) == 0 else np.int16
    lens = [int(v) for v in rng.integers(1, 60000, size=int(rng.integers(1, 7)))]
    if rng.integers(2):
        lens.insert(int(rng.integers(len(lens) + 1)), int(rng.integers(700000, 2500000)))       # one long unit
    tracks = []
    for i, n in enumerate(lens):
        x = make_track(13000 + 50 * seed + i, num_samples=n)
        tracks.append((x / 32767.).astype(np.float32) if dtype == np.float32 else x)
    plan, _ = _plan_for(dtype, N, hop, int(rng.choice([6, 12, 24])), bool(rng.integers(2)),
                        ratio=float(rng.choice([0.25, 0.5, 0.75])), max_bins=[None, 3][int(rng.integers(2))])
    dev, T = _run_all(plan, tracks, hop)
    units = _lib.make_units(lens, T)
    rows = sum(T)
    monkeypatch.setenv("MMB_HOST_GROUP_MB", str(int(rng.choice([1, 2, 8, 64]))))
    names = [k for k in ("log", "diff", "odf", "spec", "filt") if rng.integers(3) or k == "diff"]
    out = {k: np.full(dev[k].shape, np.nan, np.float32) for k in names}
    plan.run_host(np.concatenate(tracks), units, out)
    for k in names:
        assert np.array_equal(out[k], dev[k]), (seed, N, hop, lens, k)


def test_silent_frames_of_packed_pairs_are_exactly_zero():
    """frame_size 1024 transforms two frames in one complex FFT; the split separates them only up to rounding, so a
    silent frame would read about 1e-7 of a full-scale partner.  Frames without a non-zero sample after windowing are
    detected and come out as exact zeros like in the reference: the first frame of an `origin='online'` framing (one
    sample under the Hann window's zero last tap), a stretch of digital silence inside a loud track (even and odd
    frame of a pair), the zero padding around a 1024-sample track.  Loud frames next to them stay within the bound."""
    from oracle import frontend as O
    from madmom_b200.synth import make_track
    loud = make_track(321, num_samples=60000).astype(np.int16)
    hole = loud.copy()
    hole[20000:26000] = 0                                       # several frames of digital silence
    tracks = [loud, hole, make_track(322, num_samples=1024), make_track(323, num_samples=3)]
    for origin in (511, 0, -512):
        plan, recipe = _plan_for(np.int16, 1024, 441., 12, False, mul=5., origin=origin)
        got, T = _run_all(plan, tracks, 441.)
        r0, silent_rows = 0, 0
        for x, t in zip(tracks, T):
            ref = O.run_recipe(x, recipe, stages=True)
            sl = {k: v[r0:r0 + t] for k, v in got.items()}
            for k in ("filt", "log", "diff", "odf"):
                assert_close(sl[k], ref[k], what="origin %d, %d samples, %s" % (origin, len(x), k))
            zero = np.flatnonzero(np.abs(ref["spec"]).max(axis=1) == 0)      # the reference's exactly silent frames
            silent_rows += len(zero)
            assert np.all(sl["filt"][zero] == 0) and np.all(sl["log"][zero] == 0), (origin, len(x), zero[:8])
            r0 += t
        assert silent_rows >= 8                                  # the cases above do occur


def test_errors():
    from madmom_b200 import _lib
    with pytest.raises(NotImplementedError):
        _lib.FrontEndPlan(512, 441., np.hanning(512))
    plan = _lib.FrontEndPlan(2048, 441., np.hanning(2048) / 32767.)
    pcm = torch.zeros(10000, dtype=torch.int16, device="cuda")
    units = _lib.make_units([10000], [23])
    with pytest.raises(NotImplementedError):
        plan.run(pcm, units, {"log": torch.zeros((23, 0), device="cuda")})
    with pytest.raises(ValueError):
        plan.run(pcm, units, {})


@pytest.mark.parametrize("seed", range(int(__import__("os").environ.get("MMB_FUZZ_SEEDS", "40"))))
def test_random_recipes_vs_oracle(seed):
    """Differential test over the parameter space the plan accepts: frame size, hop (integer and fractional), bands
    per octave, filter normalisation, log scaling, lag ratio, frequency max-filter, signed / rectified differences,
    circular shift, origin, chunk size, int16 / float32 samples, on a ragged batch of random track lengths (shorter
    than a frame, shorter than a hop, one sample).  Every stage of every track against the oracle.
    Bounds: raw bins one float32 rounding; features 1e-5 + 1e-4 |ref| for mul / add <= 10 (the reference's recipes use
    1 and 5: the logarithm multiplies the FP32 transform's absolute noise floor by mul / (y ln 10)).
    40 seeds in the suite; MMB_FUZZ_SEEDS=400 was run in round 2.  Its first run found that N = 1024, which packs two
    frames into one complex FFT, read 1e-5 instead of 0 in frames that are exactly silent after windowing (rounding
    noise of the partner frame); the kernel now detects such frames and writes exact zeros, and all 400 pass the
    plain bound for every frame size."""
    from oracle import frontend as O
    from madmom_b200.synth import make_track
    rng = np.random.default_rng(7000 + seed)
    N = int(rng.choice([1024, 2048, 4096]))
    hop = float(rng.choice([441., 220.5, 882., 512., 147., 300.25, 1000.]))
    bpo = int(rng.choice([3, 6, 12, 24]))
    norm = bool(rng.integers(2))
    mul = float(rng.choice([1., 5., 0.5, 10.]))
    add = float(rng.choice([1., 1., 2.5]))
    ratio = float(rng.choice([0.25, 0.5, 0.75]))
    max_bins = [None, 3, 3, 5][int(rng.integers(4))]
    positive = bool(rng.integers(4) != 0)
    circ = bool(rng.integers(4) == 0)
    origin = int(rng.choice([0, 0, 0, N // 2 - 1, -N // 2, 37, -250]))
    chunk_frames = int(rng.choice([0, 0, 7, 13, 64]))
    dtype = np.float32 if rng.integers(3) == 0 else np.int16
    lens = [int(v) for v in rng.integers(1, 40000, size=5)] + [1, int(hop) - 1 or 1, N - 1, N, N + 1]
    rng.shuffle(lens)
    tracks = []
    for i, n in enumerate(lens):
        x = make_track(9000 + 100 * seed + i, num_samples=n)
        tracks.append((x / 32767.).astype(np.float32) if dtype == np.float32 else x)
    plan, recipe = _plan_for(dtype, N, hop, bpo, norm, mul=mul, add=add, ratio=ratio, max_bins=max_bins,
                             positive=positive, circ=circ, origin=origin, chunk_frames=chunk_frames)
    got, T = _run_all(plan, tracks, hop)
    what = "seed %d: N=%d hop=%g bpo=%d norm=%s mul=%g add=%g ratio=%g max_bins=%s pos=%s circ=%s origin=%d chunk=%d %s" % (
        seed, N, hop, bpo, norm, mul, add, ratio, max_bins, positive, circ, origin, chunk_frames, np.dtype(dtype).name)
    r0 = 0
    for x, t in zip(tracks, T):
        assert t == O.num_frames(len(x), hop), what
        ref = O.run_recipe(x, recipe, stages=True)
        sl = {k: v[r0:r0 + t] for k, v in got.items()}
        w = what + " | track of %d samples" % len(x)
        assert_bins_close(sl["stft"], ref["stft"], w + " stft")
        assert_bins_close(sl["spec"], ref["spec"], w + " spec")
        assert_close(sl["filt"], ref["filt"], what=w + " filt")
        assert_close(sl["log"], ref["log"], what=w + " log")
        assert_close(sl["diff"], ref["diff"], what=w + " diff")
        assert_close(sl["odf"], ref["odf"], what=w + " odf")
        r0 += t
