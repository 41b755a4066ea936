"""mmb_peak_picking_static: the entry point that allocates nothing (flags in a static device array of the library).
Collected last on purpose.  Called through the C ABI (ctypes binding in
madmom_b200._lib) and compared with the existing entry point mmb_peak_picking: frame indices are integers, so the
comparison is bit-exact, on two calls in a row with different inputs (the second call finds the first call's flags
in the shared array and must not be affected by them)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _activations(seed, lens):
    rng = np.random.default_rng(seed)
    parts = [(rng.random(n).astype(np.float32) ** 4) * 3 for n in lens]
    return np.concatenate(parts) if parts else np.zeros(0, np.float32)


def _reference_entry_point(act, seg, args):
    """mmb_peak_picking (the existing entry point): list of per-segment index arrays."""
    from madmom_b200 import _lib
    peaks, counts = _lib.peak_picking(act, *args, seg_offsets=seg)
    peaks, counts = peaks.cpu().numpy(), counts.cpu().numpy()
    offs = seg.cpu().numpy() if seg is not None else np.array([0, act.shape[0]])
    return [peaks[int(offs[s]):int(offs[s]) + int(counts[s])] for s in range(len(counts))]


def test_static_peak_picking_matches_on_two_calls_in_a_row():
    from madmom_b200 import _lib
    dev = torch.device("cuda:0")
    cap = 20000
    # buffers of both calls are made up front: nothing is allocated between / inside the calls under test
    cases = []
    for seed, lens, args in ((1, [561, 0, 40000, 17, 1, 3000], (0.5, 0, 30, 0, 2, 10)),
                             (2, [3000, 123], (0.1, 5, 3, 7, 0, 0)),          # Hamming smoothing, float64 path
                             (3, [100000], (1.1, 0, 0, 0, 1, 1))):
        x = torch.from_numpy(_activations(seed, lens)).to(dev)
        seg = torch.from_numpy(np.concatenate(([0], np.cumsum(lens))).astype(np.int64)).to(dev)
        kern = torch.from_numpy(np.hamming(args[1])).to(dev) if args[1] > 1 else None
        peaks = torch.full((len(lens), cap), -1, dtype=torch.int32, device=dev)
        counts = torch.full((len(lens),), -1, dtype=torch.int32, device=dev)
        cases.append((x, seg, args, kern, peaks, counts))
    want = [_reference_entry_point(x, seg, args) for x, seg, args, _, _, _ in cases]

    def run(case):
        x, seg, (thr, smooth, pre_avg, post_avg, pre_max, post_max), kern, peaks, counts = case
        _lib.peak_picking_static(x, thr, peaks, counts, smooth=smooth, smooth_kernel=kern, pre_avg=pre_avg,
                                 post_avg=post_avg, pre_max=pre_max, post_max=post_max, seg_offsets=seg)

    run(cases[0])                                   # first call on this device: looks the array's address up
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(dev)[0]
    reserved0 = torch.cuda.memory_reserved(dev)
    for case in cases + cases[::-1]:                # calls in a row, every one after another input's flags
        run(case)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] >= free0 and torch.cuda.memory_reserved(dev) == reserved0
    for (x, seg, args, kern, peaks, counts), ref in zip(cases, want):
        got_counts = counts.cpu().numpy()
        got = peaks.cpu().numpy()
        assert list(got_counts) == [len(r) for r in ref]
        for s, r in enumerate(ref):
            assert len(r) <= cap
            assert np.array_equal(got[s, :len(r)], r), (args, s)
            assert (got[s, len(r):] == -1).all()                  # nothing written past the peaks found
    assert sum(len(r) for r in want[0]) > 100                      # the comparison is not vacuous


def test_static_peak_picking_capacity_and_limits():
    from madmom_b200 import _lib
    dev = torch.device("cuda:0")
    x = torch.from_numpy(_activations(4, [20000])).to(dev)
    ref = _reference_entry_point(x, None, (0.1, 0, 0, 0, 1, 1))[0]
    assert len(ref) > 64
    peaks = torch.full((1, 64), -1, dtype=torch.int32, device=dev)
    counts = torch.zeros((1,), dtype=torch.int32, device=dev)
    _lib.peak_picking_static(x, 0.1, peaks, counts, pre_max=1, post_max=1)
    assert int(counts[0]) == len(ref)                              # found, not stored: the caller sees the overflow
    assert np.array_equal(peaks.cpu().numpy()[0], ref[:64])
    # an empty input: one launch, count 0
    empty = torch.zeros((0,), dtype=torch.float32, device=dev)
    _lib.peak_picking_static(empty, 0.5, peaks, counts)
    assert int(counts[0]) == 0
    # more frames than the static array holds: refused with a message, nothing launched
    lib = _lib.load()
    rc = lib.mmb_peak_picking_static(x.data_ptr(), None, 1, _lib.STATIC_PEAK_MAX_FRAMES + 1, 0.5, 0, None, 0, 0, 1, 1,
                                     peaks.data_ptr(), 64, counts.data_ptr(), None)
    assert rc == _lib.MMB_E_INVALID
    assert b"MMB_STATIC_PEAK_MAX_FRAMES" in lib.mmb_last_error()
    with pytest.raises(ValueError):
        _lib.peak_picking_static(x, 0.5, peaks[:, :0].contiguous(), counts)      # zero capacity
    torch.cuda.synchronize()
