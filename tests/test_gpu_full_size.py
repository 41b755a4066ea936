"""BASELINE.json configurations at their full size (1024 tracks x 30 s per GPU), checked through properties that do
not need the CPU oracle on all of it: per-track results are independent of batching, chunking and launch, the flux is
the row sum of the difference, lead-in rows are zero, runs are bit-reproducible; a few tracks are also compared with
the oracle."""
import numpy as np
import pytest

from conftest import assert_close

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

TRACKS, SECONDS = 1024, 30.0


@pytest.fixture(scope="module")
def batch():
    from madmom_b200.synth import make_batch_torch
    pcm = make_batch_torch(TRACKS, SECONDS, device="cuda", seed=1000)
    return pcm, [pcm.shape[1]] * TRACKS


def test_config2_logfiltspecflux_full_size(batch):
    from oracle import frontend as O
    from madmom_b200 import recipe as R
    from madmom_b200.batch import BatchFrontEnd
    pcm, lengths = batch
    fe = BatchFrontEnd(R.LOGFILTSPECFLUX)
    out = fe.run_device(pcm, lengths, ("diff", "odf"))
    diff, odf = out["diff"], out["odf"]
    T = R.num_frames(lengths[0], 441.)
    assert T == 3000 and diff.shape == (TRACKS * T, 81) and odf.shape == (TRACKS * T,)
    assert bool(torch.isfinite(diff).all()) and float(diff.min()) >= 0.0
    d = fe.plans[0].diff_frames
    lead = diff.view(TRACKS, T, 81)[:, :d]
    assert not bool(lead.any())                                   # first d rows of every track (spectrogram.py:984-989)
    # flux = row sum (features/onsets.py:163-190); the kernel adds 3 values per lane and then a shuffle tree
    ref_sum = diff.double().sum(dim=1)
    assert float((odf.double() - ref_sum).abs().max()) <= 1e-5 * max(1.0, float(ref_sum.max()))
    # bit-reproducible
    again = fe.run_device(pcm, lengths, ("diff", "odf"))
    assert torch.equal(again["diff"], diff) and torch.equal(again["odf"], odf)
    # independent of batching: tracks on their own, and a different chunk size, give the same bits
    rng = np.random.default_rng(0)
    picks = sorted(set([0, TRACKS - 1] + list(rng.integers(0, TRACKS, 6))))
    small = BatchFrontEnd(R.LOGFILTSPECFLUX, chunk_frames=97)
    for i in picks:
        alone = fe.run_device(pcm[i], [lengths[i]], ("diff",))["diff"]
        assert torch.equal(alone, diff[i * T:(i + 1) * T]), i
        assert torch.equal(small.run_device(pcm[i], [lengths[i]], ("diff",))["diff"], alone), i
    # against the oracle (two tracks)
    for i in picks[:2]:
        x = pcm[i].cpu().numpy()
        st = O.run_recipe(x, dict(frame_size=2048, hop_size=441., num_bands=12, norm_filters=False, mul=1, add=1,
                                  diff_ratio=0.5, positive_diffs=True), stages=True)
        assert_close(diff[i * T:(i + 1) * T].cpu().numpy(), st["diff"], what="config 2 diff, track %d" % i)
        assert_close(odf[i * T:(i + 1) * T].cpu().numpy(), st["odf"], what="config 2 flux, track %d" % i)


def test_config3_multires_full_size(batch):
    from oracle import frontend as O
    from madmom_b200 import recipe as R
    from madmom_b200.batch import BatchFrontEnd
    pcm, lengths = batch
    fe = BatchFrontEnd(R.RNN_BEAT)
    stack = fe.run_device(pcm, lengths, ("stack",))["stack"]
    T = 3000
    assert stack.shape == (TRACKS * T, 266) and bool(torch.isfinite(stack).all())
    # column blocks: [log 39 | diff 39 | log 45 | diff 45 | log 49 | diff 49]; each block equals the single-recipe run
    c = 0
    for rec, plan in zip(R.RNN_BEAT, fe.plans):
        b = plan.num_bands
        one = BatchFrontEnd([rec]).run_device(pcm[:64], lengths[:64], ("log", "diff"))
        assert torch.equal(one["log"], stack[:64 * T, c:c + b])
        assert torch.equal(one["diff"], stack[:64 * T, c + b:c + 2 * b])
        c += 2 * b
    assert c == 266
    picks = [3, 700]
    for i in picks:
        alone = fe.run_device(pcm[i], [lengths[i]], ("stack",))["stack"]
        assert torch.equal(alone, stack[i * T:(i + 1) * T]), i
        # against the oracle: the (3000, 266) stack of this track, all three frame sizes
        x = pcm[i].cpu().numpy()
        cols = []
        for N in (1024, 2048, 4096):
            st = O.run_recipe(x, dict(frame_size=N, hop_size=441., num_bands=6, norm_filters=True, mul=1, add=1,
                                      diff_ratio=0.5, positive_diffs=True), stages=True)
            cols += [st["log"], st["diff"]]
        assert_close(stack[i * T:(i + 1) * T].cpu().numpy(), np.hstack(cols), what="config 3 stack, track %d" % i)


def test_config4_ten_hour_signal_shards():
    """One 10-hour signal (1 587 600 000 samples): the frame ranges of 8 ranks, run one after the other on this GPU
    and concatenated like the all-gather does, are bit-identical to the single launch over the whole signal; shard
    boundaries are additionally compared with the oracle on their PCM neighbourhood."""
    from oracle import frontend as O
    from madmom_b200 import parallel, recipe as R
    from madmom_b200.batch import BatchFrontEnd
    from madmom_b200.synth import make_batch_torch
    n = 10 * 3600 * 44100
    x = make_batch_torch(1, n / 44100.0, device="cuda", seed=7).reshape(-1)
    assert x.numel() == n
    rec = R.SUPERFLUX
    whole = BatchFrontEnd(rec).run_device(x, [n], ("diff",))["diff"]
    T = R.num_frames(n, rec.hop_size)
    assert whole.shape == (T, 140) and T == 3600000
    fe = parallel.LongSignalFrontEnd(rec)
    world = 8
    r0 = 0
    for rank in range(world):
        sh = fe.shard(n, world, rank)
        local = fe.run_local(x[sh.sample_lo:sh.sample_hi].contiguous(), sh, ("diff",))["diff"]
        rows = min(local.shape[0], T - r0)
        assert torch.equal(local[:rows], whole[r0:r0 + rows]), rank
        r0 += rows
    assert r0 == T
    # oracle on the neighbourhood of two shard boundaries (frames [f - 3, f + 3))
    d = fe.plan.diff_frames
    for f in (T // 8, 5 * (T // 8)):
        lo = max(O.frame_start(f - 3 - d, 2048, 441.), 0)
        hi = O.frame_start(f + 3, 2048, 441.) + 2048
        seg = x[lo:hi].cpu().numpy()
        # frames f-3-d .. f+2 cut out of the segment with their global start samples
        fr = np.stack([seg[O.frame_start(g, 2048, 441.) - lo:O.frame_start(g, 2048, 441.) - lo + 2048]
                       for g in range(f - 3 - d, f + 3)])
        win = O.fft_window(np.hanning, 2048, np.int16)
        spec = O.spec(O.stft(fr, win))
        fb = O.logarithmic_filterbank(O.fft_frequencies(1024, 44100), num_bands=24, norm_filters=False)
        logf = O.logarithmic(O.filtered(spec, fb))
        want = O.difference(logf, d, 3, True)[d:]
        assert_close(whole[f - 3:f + 3].cpu().numpy(), want, what="config 4 boundary %d" % f)


def test_config5_complexflux_full_size(batch):
    from oracle import frontend as O
    from madmom_b200 import recipe as R
    from madmom_b200.batch import BatchFrontEnd, complex_flux_batch
    pcm, lengths = batch
    fe = BatchFrontEnd([R.COMPLEXFLUX])
    odf = complex_flux_batch(fe, pcm, lengths)
    T = R.num_frames(lengths[0], R.COMPLEXFLUX.hop_size)
    assert T == 6000 and odf.shape == (TRACKS * T,) and bool(torch.isfinite(odf).all())
    assert torch.equal(complex_flux_batch(fe, pcm, lengths), odf)                     # reproducible
    # independent of the grouping of tracks into scratch-sized launches
    assert torch.equal(complex_flux_batch(fe, pcm, lengths, scratch_bytes=300 << 20), odf)
    for i in (0, 517):
        alone = complex_flux_batch(fe, pcm[i], [lengths[i]])
        assert torch.equal(alone, odf[i * T:(i + 1) * T]), i
    i = 517
    st = O.run_recipe(pcm[i].cpu().numpy(), dict(O.SUPERFLUX_200, circular_shift=True), stages=True)
    want = O.complex_flux(st["log"], st["stft"], st["filterbank"], st["diff_frames"])
    assert_close(odf[i * T:(i + 1) * T].cpu().numpy(), want, what="config 5 ODF")


def test_synthetic_generator_is_reproducible():
    """bench.py's long-signal check and the multi-rank runs draw 'the same' signal from the same seed: the draw must
    not depend on the order of the atomic adds of bursts that overrun the end of a track (0.6 s tracks: most of the
    0.5 s bursts overrun)."""
    import torch
    from madmom_b200.synth import make_batch_torch
    a = make_batch_torch(64, 0.6, device="cuda", seed=11)
    for _ in range(3):
        assert torch.equal(a, make_batch_torch(64, 0.6, device="cuda", seed=11))
    assert int(a[:, -1].abs().max()) < 32767 and a.float().std() > 100
