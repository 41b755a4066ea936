"""Onset detection functions on top of the fused front end, and peak picking on the device.

Mirror of madmom/features/onsets.py: `spectral_flux` (:163-190), `superflux` (:193-240), the other spectral
onset detection functions (:102-160, :320-581), `SpectralOnsetProcessor`
(:584-720), `peak_picking` (:875-963), `OnsetPeakPickingProcessor` (:966-1180, offline mode) and
`madmom.utils.combine_events` (madmom/utils/__init__.py:275-330).

The flux reductions are the row-sum epilogue of the fused kernel; peak picking runs as two small CUDA kernels
(`mmb_peak_picking`), so an onset pipeline copies only the final handful of frame indices back to the host.
"""
import inspect

import numpy as np

from ..processors import Processor, SequentialProcessor
from ..audio.stft import DeviceResult
from ..audio.spectrogram import Spectrogram, SpectrogramDifference


def _as_spectrogram(spectrogram):
    if not isinstance(spectrogram, Spectrogram):
        spectrogram = Spectrogram(spectrogram)
    return spectrogram


def spectral_flux(spectrogram, diff_frames=None):
    """Sum of the positive lagged differences per frame (madmom/features/onsets.py:163-190)."""
    diff = SpectrogramDifference(_as_spectrogram(spectrogram), diff_frames=diff_frames, positive_diffs=True)
    return diff.flux()


def superflux(spectrogram, diff_frames=None, diff_max_bins=3):
    """SuperFlux: spectral flux against a frequency-maximum-filtered lagged frame
    (madmom/features/onsets.py:193-240)."""
    diff = SpectrogramDifference(_as_spectrogram(spectrogram), diff_frames=diff_frames,
                                 diff_max_bins=diff_max_bins, positive_diffs=True)
    return diff.flux()


class _ComplexFlux(DeviceResult):
    """ODF of complex_flux: the fused kernel writes |local group delay| / pi per bin and the SuperFlux difference;
    `mmb_complex_flux` applies the temporal maximum filter, the per-band minimum and the weighted sum."""

    def __init__(self, diff, temporal_filter):
        DeviceResult.__init__(self, diff._chain.copy())
        self.diff = diff
        self.temporal_filter = int(temporal_filter)

    def _shape(self):
        return (self.diff.shape[0],)

    def _compute(self):
        import torch
        from .. import _lib
        out = self._chain.run(("lgd", "diff"))
        starts, lens, _ = self.diff.filterbank.band_table()
        num_bins = self.diff.filterbank.shape[0]
        lo = np.maximum(starts.astype(np.int64) - 1, 0).astype(np.int32)
        hi = np.minimum(starts.astype(np.int64) + lens + 1, num_bins).astype(np.int32)
        dev = out["lgd"].device
        return _lib.complex_flux(out["lgd"], out["diff"], torch.from_numpy(lo).to(dev), torch.from_numpy(hi).to(dev),
                                 self.temporal_filter)


def complex_flux(spectrogram, diff_frames=None, diff_max_bins=3, temporal_filter=3, temporal_origin=0):
    """ComplexFlux: SuperFlux weighted by a local-group-delay mask (madmom/features/onsets.py:245-317).
    The spectrogram must be filtered and its STFT circularly shifted, as in the reference's ComplexFlux program."""
    spectrogram = _as_spectrogram(spectrogram)
    if temporal_origin != 0:
        raise NotImplementedError("temporal_origin != 0 is not available on the GPU path")
    if temporal_filter > 1 and temporal_filter % 2 == 0:
        raise NotImplementedError("even temporal_filter sizes are not available on the GPU path")
    if spectrogram._chain.filterbank is None or getattr(spectrogram, "_generic", False):
        raise NotImplementedError("complex_flux needs a filtered (optionally logarithmic) spectrogram in the "
                                  "reference's canonical order on the GPU path")
    if spectrogram.stft.fft_size != 2048:
        raise NotImplementedError("complex_flux is implemented for frame_size 2048 only")
    diff = SpectrogramDifference(spectrogram, diff_frames=diff_frames, diff_max_bins=diff_max_bins,
                                 positive_diffs=True)
    return _ComplexFlux(diff, temporal_filter)


EPSILON = np.spacing(1)


class _SpectralODF(DeviceResult):
    """One of the bin-wise onset detection functions of madmom/features/onsets.py:102-160, 320-581, computed by
    `mmb_spectral_odf` (one warp per frame) from the spectrogram or the STFT on the device."""

    def __init__(self, spectrogram, mode, diff_frames=1, epsilon=EPSILON):
        DeviceResult.__init__(self, spectrogram._chain.copy())
        self.spectrogram = spectrogram
        self.mode = mode
        self.diff_frames = int(diff_frames)
        self.epsilon = float(epsilon)

    def _shape(self):
        return (self.spectrogram.shape[0],)

    def _compute(self):
        from .. import _lib
        if _lib.ODF_MODES[self.mode] <= 3:
            return _lib.spectral_odf(self.mode, spec=self.spectrogram.tensor, diff_frames=self.diff_frames,
                                     epsilon=self.epsilon)
        return _lib.spectral_odf(self.mode, stft=self.spectrogram.stft.tensor, diff_frames=self.diff_frames,
                                 epsilon=self.epsilon)


def _default_diff_frames(spectrogram, diff_frames):
    if diff_frames is not None:
        return int(diff_frames)
    from ..recipe import diff_frames as _diff_frames
    from ..audio.spectrogram import DIFF_RATIO
    return _diff_frames(DIFF_RATIO, hop_size=spectrogram.stft.frames.hop_size,
                        frame_size=spectrogram.stft.frames.frame_size, window=spectrogram.stft.window)


def _unfiltered(spectrogram):
    # the reference compares np.shape(phase) with np.shape(spectrogram) (onsets.py:438-440, 503-505)
    if tuple(spectrogram.shape) != tuple(spectrogram.stft.shape):
        raise ValueError("spectrogram and phase must be of same shape")


def high_frequency_content(spectrogram):
    """Mean over the bins of magnitude times bin number (madmom/features/onsets.py:102-127).  float32 on the
    device (the reference's result is float64)."""
    return _SpectralODF(_as_spectrogram(spectrogram), "high_frequency_content")


def spectral_diff(spectrogram, diff_frames=None):
    """Sum of the squared positive differences (madmom/features/onsets.py:130-160)."""
    spectrogram = _as_spectrogram(spectrogram)
    if isinstance(spectrogram, SpectrogramDifference):
        raise NotImplementedError("spectral_diff of a SpectrogramDifference is not available on the GPU path; "
                                  "pass the spectrogram itself")
    return _SpectralODF(spectrogram, "spectral_diff", _default_diff_frames(spectrogram, diff_frames))


def modified_kullback_leibler(spectrogram, diff_frames=1, epsilon=EPSILON):
    """Modified Kullback-Leibler (madmom/features/onsets.py:320-362)."""
    if epsilon <= 0:
        raise ValueError("a positive value must be added before division")
    return _SpectralODF(_as_spectrogram(spectrogram), "modified_kullback_leibler", diff_frames, epsilon)


def phase_deviation(spectrogram):
    """Mean absolute second difference of the phase (madmom/features/onsets.py:390-414)."""
    return _SpectralODF(_as_spectrogram(spectrogram), "phase_deviation")


def weighted_phase_deviation(spectrogram):
    """Phase deviation weighted by the magnitudes (madmom/features/onsets.py:417-446)."""
    spectrogram = _as_spectrogram(spectrogram)
    _unfiltered(spectrogram)
    return _SpectralODF(spectrogram, "weighted_phase_deviation")


def normalized_weighted_phase_deviation(spectrogram, epsilon=EPSILON):
    """Weighted phase deviation over the mean magnitude (madmom/features/onsets.py:449-479)."""
    if epsilon <= 0:
        raise ValueError("a positive value must be added before division")
    spectrogram = _as_spectrogram(spectrogram)
    _unfiltered(spectrogram)
    return _SpectralODF(spectrogram, "normalized_weighted_phase_deviation", epsilon=epsilon)


def complex_domain(spectrogram):
    """Complex domain (madmom/features/onsets.py:482-549)."""
    spectrogram = _as_spectrogram(spectrogram)
    _unfiltered(spectrogram)
    return _SpectralODF(spectrogram, "complex_domain")


def rectified_complex_domain(spectrogram, diff_frames=None):
    """Complex domain restricted to bins with rising magnitude (madmom/features/onsets.py:552-581)."""
    spectrogram = _as_spectrogram(spectrogram)
    _unfiltered(spectrogram)
    return _SpectralODF(spectrogram, "rectified_complex_domain", _default_diff_frames(spectrogram, diff_frames))


class SpectralOnsetProcessor(SequentialProcessor):
    """Signal -> frames -> STFT -> spectrogram [-> filterbank] [-> log] -> onset detection function
    (madmom/features/onsets.py:584-683).  The whole chain is one kernel launch."""
    METHODS = ["superflux", "complex_flux", "high_frequency_content", "spectral_diff", "spectral_flux",
               "modified_kullback_leibler", "phase_deviation", "weighted_phase_deviation",
               "normalized_weighted_phase_deviation", "complex_domain", "rectified_complex_domain"]

    def __init__(self, onset_method="spectral_flux", **kwargs):
        from ..audio.signal import SignalProcessor, FramedSignalProcessor
        from ..audio.stft import ShortTimeFourierTransformProcessor
        from ..audio.spectrogram import (SpectrogramProcessor, FilteredSpectrogramProcessor,
                                         LogarithmicSpectrogramProcessor)
        if not inspect.isfunction(onset_method) and any(odf in onset_method for odf in ("phase", "complex")):
            kwargs["circular_shift"] = True
        kwargs["num_channels"] = 1
        processors = [SignalProcessor(**kwargs), FramedSignalProcessor(**kwargs),
                      ShortTimeFourierTransformProcessor(**kwargs), SpectrogramProcessor(**kwargs)]
        if kwargs.get("filterbank") is not None:
            processors.append(FilteredSpectrogramProcessor(**kwargs))
        if kwargs.get("log") is not None:
            processors.append(LogarithmicSpectrogramProcessor(**kwargs))
        if not inspect.isfunction(onset_method):
            try:
                onset_method = globals()[onset_method]
            except KeyError:
                raise ValueError("%s not a valid onset detection function, choose %s." % (onset_method,
                                                                                           self.METHODS))
        processors.append(onset_method)
        super(SpectralOnsetProcessor, self).__init__(processors)


# ------------------------------------------------------------------------------------------------ peak picking
def _device_vector(activations):
    """1-D float32 CUDA tensor of an activation function (device results stay where they are)."""
    import torch
    from .._engine import current_device
    if isinstance(activations, DeviceResult):
        t = activations.tensor
    elif isinstance(activations, torch.Tensor):
        t = activations
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(activations)))
    if t.dim() != 1:
        raise ValueError("`activations` must be a 1D array on the GPU path")       # the 2D case is not offered
    return t.to(device="cuda:%d" % current_device(), dtype=torch.float32).contiguous()


def peak_picking(activations, threshold, smooth=None, pre_avg=0, post_avg=0, pre_max=1, post_max=1):
    """Frame indices of the peaks of a 1-D activation function (madmom/features/onsets.py:875-963).
    All lengths are in frames.  Returns a numpy int array like np.nonzero()[0] does in the reference."""
    from .. import _lib
    if smooth is None:
        smooth = 0
    if isinstance(smooth, np.ndarray):
        raise NotImplementedError("explicit smoothing kernels are not available on the GPU path")
    if smooth == 1 or smooth < 0:
        raise ValueError("can't create a smoothing kernel of size %d" % smooth)
    act = _device_vector(activations)
    peaks, counts = _lib.peak_picking(act, threshold, int(smooth), int(pre_avg), int(post_avg), int(pre_max),
                                      int(post_max))
    n = int(counts[0].item())
    return peaks[:n].cpu().numpy().astype(np.int64)


def combine_events(events, delta, combine="mean"):
    """Merge events closer than `delta` (madmom/utils/__init__.py:275-330); a handful of scalars on the host."""
    delta += 1e-12
    if len(events) <= 1:
        return events
    events = np.array(events, dtype=float)
    if events.ndim > 1:
        raise ValueError("only 1-dimensional events supported.")
    kept = 0
    left = events[0]
    for right in events[1:]:
        if right - left <= delta:
            if combine == "mean":
                left = events[kept] = 0.5 * (right + left)
            elif combine == "left":
                left = events[kept] = left
            elif combine == "right":
                left = events[kept] = right
            else:
                raise ValueError("don't know how to combine two events with %s" % combine)
        else:
            kept += 1
            left = events[kept] = right
    return events[:kept + 1]


class OnsetPeakPickingProcessor(Processor):
    """Offline peak picking of an onset activation function; times in seconds
    (madmom/features/onsets.py:966-1107).  Online (frame-by-frame) mode is not available on the GPU path."""
    FPS = 100
    THRESHOLD = 0.5
    SMOOTH = 0.
    PRE_AVG = 0.
    POST_AVG = 0.
    PRE_MAX = 0.
    POST_MAX = 0.
    COMBINE = 0.03
    DELAY = 0.

    def __init__(self, threshold=THRESHOLD, smooth=SMOOTH, pre_avg=PRE_AVG, post_avg=POST_AVG, pre_max=PRE_MAX,
                 post_max=POST_MAX, combine=COMBINE, delay=DELAY, online=False, fps=FPS, **kwargs):
        if online:
            raise NotImplementedError("online peak picking is not available on the GPU path")
        self.threshold = threshold
        self.smooth = smooth
        self.pre_avg = pre_avg
        self.post_avg = post_avg
        self.pre_max = pre_max
        self.post_max = post_max
        self.combine = combine
        self.delay = delay
        self.fps = fps

    def process(self, activations, **kwargs):
        timings = np.array([self.smooth, self.pre_avg, self.post_avg, self.pre_max, self.post_max]) * self.fps
        timings = np.round(timings).astype(int)
        onsets = peak_picking(activations, self.threshold, *timings)
        onsets = onsets.astype(float) / self.fps
        if self.delay:
            onsets += self.delay
        if self.combine:
            onsets = combine_events(onsets, self.combine, "left")
        return np.asarray(onsets)

    process_offline = process
