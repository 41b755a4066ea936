"""Glue between the lazy result objects (audio/stft.py, audio/spectrogram.py) and the CUDA library: turns the
description of a chain (frames -> window/FFT -> filterbank -> log -> difference) into one cached `FrontEndPlan`
and one fused kernel launch.  Nothing here computes on the host."""
import hashlib
import threading

import numpy as np

from . import _lib
from .recipe import fft_window as _fft_window

_plans = {}
_lock = threading.Lock()


def current_device():
    import torch
    if not torch.cuda.is_available():
        raise _lib.LibraryError("madmom_b200 needs a CUDA device: the spectral front end has no CPU path")
    return torch.cuda.current_device()


def log_mode_of(log, mul, add):
    if log is None:
        return _lib.LOG_MULADD if (mul is not None or add is not None) else _lib.LOG_NONE
    for fn, mode in ((np.log10, _lib.LOG_LOG10), (np.log, _lib.LOG_LN), (np.log2, _lib.LOG_LOG2)):
        if log is fn:
            return mode
    raise NotImplementedError("log function %r is not available on the GPU path (np.log10, np.log, np.log2 or "
                              "None)" % (log,))


def _filterbank_key(fb):
    """Plan-cache key of a filterbank: shape + digest of its float32 bytes.  The digest is kept on the array (the
    processors hand the same Filterbank object to every call, like the reference's cache in
    FilteredSpectrogramProcessor) and the array is frozen from then on, so a later in-place change raises instead
    of silently running with the plan of the old weights."""
    if fb is None:
        return None
    key = getattr(fb, "_mmb_plan_key", None)
    if key is not None and not fb.flags.writeable:
        return key
    arr = np.ascontiguousarray(fb, dtype=np.float32)
    key = (arr.shape, hashlib.blake2b(arr.tobytes(), digest_size=16).digest())
    if isinstance(fb, np.ndarray):
        try:
            fb._mmb_plan_key = key
            fb.setflags(write=False)
        except (AttributeError, ValueError):
            pass
    return key


class Chain(object):
    """Everything that defines the numbers of one chain; hashable into a plan key."""

    def __init__(self, frames, window=np.hanning, circular_shift=False, fft_window=None, include_nyquist=False):
        self.frames = frames
        self.window = window
        self.circular_shift = bool(circular_shift)
        self.include_nyquist = bool(include_nyquist)
        self.fft_window = fft_window            # float64 array actually multiplied onto the samples
        self.filterbank = None
        self.log_mode, self.mul, self.add = _lib.LOG_NONE, 1.0, 0.0
        self.diff_frames, self.diff_max_bins, self.positive_diffs = 0, 0, False

    def copy(self):
        c = Chain.__new__(Chain)
        c.__dict__.update(self.__dict__)
        return c

    def resolved_fft_window(self):
        if self.fft_window is None:
            self.fft_window = _fft_window(self.window, self.frames.frame_size, self.frames.signal.dtype)
        return self.fft_window

    def plan(self, device):
        fr = self.frames
        dtype = np.dtype(fr.signal.dtype)
        fw = np.ascontiguousarray(self.resolved_fft_window(), dtype=np.float64)
        fb = self.filterbank
        key = (device, fr.frame_size, fr.hop_size, dtype.str, self.circular_shift, fw.tobytes(),
               _filterbank_key(fb), self.log_mode, float(self.mul),
               float(self.add), self.diff_frames, self.diff_max_bins, self.positive_diffs, self.include_nyquist)
        with _lock:
            plan = _plans.get(key)
            if plan is None:
                if len(_plans) > 64:
                    _plans.clear()
                # the plan's origin is 0: the FramedSignal's origin travels in the unit's sample_base
                plan = _lib.FrontEndPlan(fr.frame_size, fr.hop_size, fw, pcm_dtype=dtype, origin=0,
                                         circular_shift=self.circular_shift, filterbank=fb, log_mode=self.log_mode,
                                         mul=self.mul, add=self.add, diff_frames=self.diff_frames,
                                         diff_max_bins=self.diff_max_bins, positive_diffs=self.positive_diffs,
                                         device=device, include_nyquist=self.include_nyquist)
                _plans[key] = plan
        return plan

    def run(self, names):
        """Launch the fused kernel for this chain; returns {name: CUDA tensor} ('stack' = hstack((log, diff)))."""
        import torch
        fr = self.frames
        if fr.signal.ndim != 1:
            raise NotImplementedError("multi-channel signals are not supported on the GPU path; down-mix with "
                                      "Signal(..., num_channels=1)")
        device = current_device()
        plan = self.plan(device)
        pcm = fr.device_pcm(device)
        T = fr.num_frames
        units = np.zeros(1, dtype=_lib.TRACK_DTYPE)
        # frame i starts at int(i * hop) - N/2 - origin: the origin is a shift of the sample axis
        units[0] = (0, len(fr.signal), fr.origin, 0, T, 0)
        dev = "cuda:%d" % device
        out, bind = {}, {}
        for n in names:
            if n == "stack":
                B = plan.num_bands
                out[n] = torch.empty((T, 2 * B), dtype=torch.float32, device=dev)
                bind["log"], bind["diff"] = out[n][:, :B], out[n][:, B:]
            elif n == "odf":
                out[n] = bind[n] = torch.empty((T,), dtype=torch.float32, device=dev)
            elif n == "stft":
                out[n] = bind[n] = torch.empty((T, plan.num_bins), dtype=torch.complex64, device=dev)
            else:
                out[n] = bind[n] = torch.empty((T, plan.width(n)), dtype=torch.float32, device=dev)
        if T > 0:
            plan.run(pcm, units, bind)
        return out
