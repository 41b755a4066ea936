"""Batched front end: many tracks (or frame ranges of one long signal) through the fused kernel in one launch
per recipe.  This is the data-parallel entry point behind the processor classes; the reference's equivalent
is one Python process per file (`process_batch`, madmom/processors.py:637-713).

    fe = BatchFrontEnd(recipe.LOGFILTSPECFLUX)                  # one plan on the current CUDA device
    out = fe.run_device(pcm_cuda, lengths, ("diff", "odf"))     # tensors stay on the GPU
    out = fe.run_host(pcm_pinned, lengths, ("diff", "odf"))     # host in, host out (H2D/kernel/D2H overlapped)

Several recipes (the multi-resolution stack of RNNOnsetProcessor / RNNBeatProcessor) write directly into the
column blocks of one `(rows, sum 2 B_i)` matrix, which is the reference's np.hstack of hstack((spec, diff)).
"""
import numpy as np

from . import _lib
from . import recipe as R


def _log_mode(rec):
    if rec.log is None:
        return _lib.LOG_MULADD if (rec.mul is not None or rec.add is not None) else _lib.LOG_NONE
    if rec.log is np.log10:
        return _lib.LOG_LOG10
    if rec.log is np.log:
        return _lib.LOG_LN
    if rec.log is np.log2:
        return _lib.LOG_LOG2
    raise NotImplementedError("log function %r is not available on the GPU path (np.log10, np.log, np.log2 "
                              "or None)" % (rec.log,))


def plan_for(rec, pcm_dtype=np.int16, device=0, chunk_frames=0):
    """FrontEndPlan of one Recipe."""
    fb = rec.build_filterbank()
    d = rec.resolved_diff_frames()
    return _lib.FrontEndPlan(rec.frame_size, rec.hop_size, R.fft_window(rec.window, rec.frame_size, pcm_dtype),
                             pcm_dtype=pcm_dtype, origin=R.resolve_origin(rec.origin, rec.frame_size),
                             circular_shift=rec.circular_shift, filterbank=fb, log_mode=_log_mode(rec),
                             mul=1.0 if rec.mul is None else rec.mul, add=0.0 if rec.add is None else rec.add,
                             diff_frames=d, diff_max_bins=rec.diff_max_bins or 0,
                             positive_diffs=rec.positive_diffs, chunk_frames=chunk_frames, device=device)


class BatchFrontEnd(object):
    def __init__(self, recipes, pcm_dtype=np.int16, device=None, chunk_frames=0):
        import torch
        if device is None:
            device = torch.cuda.current_device()
        self.device = int(device)
        self.recipes = list(recipes) if isinstance(recipes, (list, tuple)) else [recipes]
        hops = {r.hop_size for r in self.recipes}
        ends = {r.end for r in self.recipes}
        if len(hops) != 1 or len(ends) != 1:
            raise ValueError("all recipes of a stack must share hop_size and end (frame counts must align)")
        self.hop_size, self.end = hops.pop(), ends.pop()
        self.pcm_dtype = np.dtype(pcm_dtype)
        self.plans = [plan_for(r, self.pcm_dtype, self.device, chunk_frames) for r in self.recipes]
        self.num_bands = [p.num_bands for p in self.plans]
        self.stack_width = 2 * sum(self.num_bands)

    # ------------------------------------------------------------------------------------------ layout
    def frame_counts(self, lengths):
        return [R.num_frames(int(n), self.hop_size, self.end) for n in lengths]

    def units(self, lengths):
        T = self.frame_counts(lengths)
        return _lib.make_units(lengths, T), T

    def width(self, name):
        if name == "stack":
            return self.stack_width
        return self.plans[0].width(name)

    def _alloc(self, names, rows, make):
        out = {}
        for n in names:
            if n == "stack":
                out[n] = make((rows, self.stack_width), "f")
            elif n == "odf":
                out[n] = make((rows,), "f")
            elif n == "stft":
                out[n] = make((rows, self.width(n)), "c")
            else:
                if len(self.plans) != 1:
                    raise ValueError("output %r needs a single recipe; a stack of recipes produces 'stack'" % n)
                out[n] = make((rows, self.width(n)), "f")
        return out

    def _per_plan_outputs(self, out):
        """[{name: tensor-or-column-view}] per plan."""
        per = [dict() for _ in self.plans]
        for n, t in out.items():
            if n == "stack":
                c = 0
                for i, b in enumerate(self.num_bands):
                    per[i]["log"] = t[:, c:c + b]
                    per[i]["diff"] = t[:, c + b:c + 2 * b]
                    c += 2 * b
            else:
                per[0][n] = t
        return per

    # ------------------------------------------------------------------------------------------ device
    def run_device(self, pcm, lengths=None, outputs=("diff",), out=None, units=None, rows=None):
        """pcm: 1-D (tracks back to back) or 2-D (equal-length tracks) CUDA tensor.  Returns {name: tensor};
        row r of every output is frame r of the concatenated batch.  Asynchronous on the current stream."""
        import torch
        if pcm.dim() == 2 and lengths is None:
            lengths = [pcm.shape[1]] * pcm.shape[0]
        pcm = pcm.reshape(-1)
        if units is None:
            units, T = self.units(lengths)
            rows = int(sum(T))
        if out is None:
            def make(shape, kind):
                return torch.empty(shape, dtype=torch.complex64 if kind == "c" else torch.float32,
                                   device="cuda:%d" % self.device)
            out = self._alloc(outputs, rows, make)
        for plan, o in zip(self.plans, self._per_plan_outputs(out)):
            if o:
                plan.run(pcm, units, o)
        return out

    # ------------------------------------------------------------------------------------------ host
    def run_host(self, pcm, lengths=None, outputs=("diff",), out=None, num_channels=1, channel=None):
        """pcm: host array / CPU tensor (pinned for full speed).  Outputs are host tensors (pinned when
        allocated here).  Returns when they are complete.  num_channels > 1: the last axis of `pcm` holds the
        interleaved channels ((tracks, samples, channels) or (total_samples, channels)); `lengths` count sample
        frames and the mix-down to mono (madmom.audio.signal.remix) happens on the device."""
        import torch
        if isinstance(pcm, np.ndarray):
            pcm = torch.from_numpy(pcm)
        if num_channels > 1:
            if pcm.shape[-1] != num_channels:
                raise ValueError("last axis of pcm must hold the %d channels" % num_channels)
            pcm = pcm.reshape(pcm.shape[:-1].numel(), num_channels) if pcm.dim() != 3 else pcm
            if pcm.dim() == 3 and lengths is None:
                lengths = [pcm.shape[1]] * pcm.shape[0]
        elif pcm.dim() == 2 and lengths is None:
            lengths = [pcm.shape[1]] * pcm.shape[0]
        pcm = pcm.reshape(-1)
        units, T = self.units(lengths)
        rows = int(sum(T))
        if out is None:
            def make(shape, kind):
                return torch.empty(shape, dtype=torch.complex64 if kind == "c" else torch.float32,
                                   pin_memory=True)
            out = self._alloc(outputs, rows, make)
        for plan, o in zip(self.plans, self._per_plan_outputs(out)):
            if o:
                plan.run_host(pcm, units, o, num_channels=num_channels, channel=channel)
        return out


    # ------------------------------------------------------------------------------------------ program
    def onsets_host(self, pcm, lengths, threshold, fps=None, smooth=0., pre_avg=0., post_avg=0., pre_max=0.,
                    post_max=0., group_tracks=128):
        """The whole onset program for a batch of tracks held in host memory (bin/LogFiltSpecFlux, bin/SuperFlux:
        SpectralOnsetProcessor -> OnsetPeakPickingProcessor.process, madmom/features/onsets.py:651-684, 1076-1107,
        without the final combine_events): PCM in, peak frame indices out.  Groups of tracks are copied to the device
        on alternating streams while the fused kernel works on the previous group; only the onset detection
        function stays on the device, peaks are picked there (mmb_peak_picking, one segment per track) and only
        the peak indices and per-track counts travel back.  Returns (peaks int32 [total], counts int32 [tracks],
        seg_offsets int64 [tracks + 1]): track i's peaks are peaks[seg[i] : seg[i] + counts[i]] (frame indices)."""
        import torch
        if len(self.plans) != 1:
            raise ValueError("onsets_host needs a single recipe")
        plan = self.plans[0]
        if isinstance(pcm, np.ndarray):
            pcm = torch.from_numpy(pcm)
        if pcm.dim() == 2 and lengths is None:
            lengths = [pcm.shape[1]] * pcm.shape[0]
        pcm = pcm.reshape(-1)
        fps = self.recipes[0].fps if fps is None else fps
        T = self.frame_counts(lengths)
        n = len(T)
        offs = np.concatenate(([0], np.cumsum(lengths))).astype(np.int64)
        seg = np.concatenate(([0], np.cumsum(T))).astype(np.int64)
        dev = torch.device("cuda", self.device)
        st = getattr(self, "_prog_state", None)
        max_group = max(int(offs[min(i + group_tracks, n)] - offs[i]) for i in range(0, n, group_tracks))
        if st is None or st["cap"] < max_group or st["rows"] < int(seg[-1]):
            st = {"cap": max_group, "rows": int(seg[-1]),
                  "slots": [torch.empty((max_group,), dtype=pcm.dtype, device=dev) for _ in range(3)],
                  "streams": [torch.cuda.Stream(device=dev) for _ in range(3)],
                  "odf": torch.empty((int(seg[-1]),), dtype=torch.float32, device=dev)}
            self._prog_state = st
        odf = st["odf"][:int(seg[-1])]
        cur = torch.cuda.current_stream(dev)
        for s in st["streams"]:
            s.wait_stream(cur)
        for g, i0 in enumerate(range(0, n, group_tracks)):
            i1 = min(i0 + group_tracks, n)
            stream, slot = st["streams"][g % 3], st["slots"][g % 3]
            a, b = int(offs[i0]), int(offs[i1])
            with torch.cuda.stream(stream):
                slot[:b - a].copy_(pcm[a:b], non_blocking=True)
                units = _lib.make_units(lengths[i0:i1], T[i0:i1])
                plan.run(slot, units, {"odf": odf[int(seg[i0]):int(seg[i1])]}, stream=stream.cuda_stream)
        for s in st["streams"]:
            cur.wait_stream(s)
        frames = np.round(np.array([smooth, pre_avg, post_avg, pre_max, post_max]) * fps).astype(int)
        peaks, counts = _lib.peak_picking(odf, threshold, int(frames[0]), int(frames[1]), int(frames[2]),
                                          int(frames[3]), int(frames[4]), torch.from_numpy(seg).to(dev))
        counts_h = counts.cpu()
        # ascending indices per segment sit at the start of the segment's range: gather the used prefixes on the device
        total = int(counts_h.sum())
        if total:
            segd = torch.from_numpy(seg[:-1]).to(dev)
            starts = torch.repeat_interleave(segd, counts.long())
            within = torch.arange(total, device=dev) - torch.repeat_interleave(
                torch.cumsum(counts.long(), 0) - counts.long(), counts.long())
            packed = peaks[starts + within].cpu()
        else:
            packed = torch.empty((0,), dtype=torch.int32)
        return packed, counts_h, seg


def complex_flux_batch(fe, pcm, lengths, temporal_filter=3, scratch_bytes=None):
    """ComplexFlux ODF (config 5 of BASELINE.json) for a batch of tracks on the device.  The per-bin local group
    delay matrix is an intermediate of rows x K floats, so the batch is walked in groups of tracks whose scratch
    fits `scratch_bytes` (default: 16 GiB or a quarter of the free device memory; large groups keep the tail of
    the persistent front-end kernel short); per group: one fused front-end launch (lgd + difference) and one
    mmb_complex_flux launch.
    `fe` is a BatchFrontEnd built from a single circular-shift recipe (recipe.COMPLEXFLUX)."""
    import torch
    plan = fe.plans[0]
    if len(fe.plans) != 1:
        raise ValueError("complex_flux_batch needs a single recipe")
    pcm = pcm.reshape(-1)
    if scratch_bytes is None:
        scratch_bytes = min(16 << 30, torch.cuda.mem_get_info(pcm.device)[0] // 4)
    T = fe.frame_counts(lengths)
    offsets = np.concatenate(([0], np.cumsum(lengths)[:-1])).astype(np.int64)
    rows_total = int(sum(T))
    dev = pcm.device
    odf = torch.empty((rows_total,), dtype=torch.float32, device=dev)
    fb = fe.recipes[0].build_filterbank()
    starts, lens, _ = fb.band_table()
    lo = torch.from_numpy(np.maximum(starts.astype(np.int64) - 1, 0).astype(np.int32)).to(dev)
    hi = torch.from_numpy(np.minimum(starts.astype(np.int64) + lens + 1, fb.shape[0]).astype(np.int32)).to(dev)
    K, B = plan.num_bins, plan.num_bands
    max_rows = max(int(scratch_bytes // (4 * (K + B))), max(T) if T else 1)
    # the scratch matrices stay with the front end between calls (no multi-GiB allocation in the steady state)
    lgd, diff = getattr(fe, "_cf_scratch", (None, None))
    if lgd is not None and (lgd.device != dev or lgd.shape[1] != K or diff.shape[1] != B):
        lgd = diff = None
    i0, row0 = 0, 0
    while i0 < len(T):
        i1, rows = i0, 0
        while i1 < len(T) and (rows + T[i1] <= max_rows or i1 == i0):
            rows += T[i1]
            i1 += 1
        if lgd is None or lgd.shape[0] < rows:
            lgd = diff = None
            fe._cf_scratch = (None, None)
            lgd = torch.empty((max(rows, 1), K), dtype=torch.float32, device=dev)
            diff = torch.empty((max(rows, 1), B), dtype=torch.float32, device=dev)
            fe._cf_scratch = (lgd, diff)
        units = _lib.make_units(lengths[i0:i1], T[i0:i1], pcm_offsets=offsets[i0:i1])
        plan.run(pcm, units, {"lgd": lgd[:rows], "diff": diff[:rows]})
        seg = torch.from_numpy(np.concatenate(([0], np.cumsum(T[i0:i1]))).astype(np.int64)).to(dev)
        odf[row0:row0 + rows] = _lib.complex_flux(lgd[:rows], diff[:rows], lo, hi, temporal_filter, seg_offsets=seg)
        i0, row0 = i1, row0 + rows
    return odf
