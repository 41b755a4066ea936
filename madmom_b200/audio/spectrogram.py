"""Spectrogram stages: magnitude, filterbank, logarithm, difference.

Mirror of madmom/audio/spectrogram.py: `Spectrogram` (:86-236), `FilteredSpectrogram` (:271-403),
`LogarithmicSpectrogram` (:479-563), `LogarithmicFilteredSpectrogram` (:664-748), `SpectrogramDifference`
(:867-1022), `SuperFluxProcessor` (:1230-1262) and the matching processors, with the reference's constructor
signatures, attribute graph (`.stft.frames`, `.filterbank`, `.bin_frequencies`, `.spectrogram`, `.diff_frames`,
`.mul`, `.add`, ...) and error behaviour.

Every class is a lazy `DeviceResult`: building the object extends the description of the chain; the first
access to the data launches ONE fused kernel for the whole chain up to that stage (framing, window, FFT,
magnitude, filterbank, log, max-filter, difference), and only that stage's matrix is written to HBM.
"""
import numpy as np

from ..processors import Processor, SequentialProcessor
from ..recipe import diff_frames as _diff_frames  # noqa: F401  (same name as the reference helper)
from .. import _engine
from .filters import Filterbank, LogarithmicFilterbank
from .stft import DeviceResult, ShortTimeFourierTransform, ShortTimeFourierTransformProcessor

FILTERBANK = LogarithmicFilterbank
NUM_BANDS = 12
FMIN = 30.
FMAX = 17000.
FREF = 440.
NORM_FILTERS = True
UNIQUE_FILTERS = True
LOG = np.log10
MUL = 1.
ADD = 1.
DIFF_RATIO = 0.5
DIFF_FRAMES = None
DIFF_MAX_BINS = None
POSITIVE_DIFFS = False


def spec(stft):
    """Magnitudes of a complex STFT, np.abs (madmom/audio/spectrogram.py:22-38).  A device STFT stays on the device
    (the `Spectrogram` class); a plain array is answered on the host like `stft.phase()` does."""
    if isinstance(stft, ShortTimeFourierTransform):
        return Spectrogram(stft)
    return np.abs(stft)


class Spectrogram(DeviceResult):
    """Magnitude spectrogram |STFT|, (num_frames, num_bins) float32.

    `_generic`: the fused kernel computes the stages in the reference's canonical order (filter, logarithm,
    difference).  The reference's classes also accept any other order -- `LogarithmicSpectrogram(...).filter()`,
    the logarithm of a logarithm, the difference of a difference (tests/test_audio_spectrogram.py:321-324) -- and
    just apply `np.dot` / `log` / the lagged difference to whatever matrix they are given.  Such results are marked
    generic: they are computed from their source's device matrix (a plain float32 matrix product, an element-wise
    logarithm, `mmb_matrix_diff`), off the fused path, and everything built on them stays generic."""
    _output = "spec"
    _generic = False

    def __init__(self, stft, **kwargs):
        self._source = None
        if isinstance(stft, Spectrogram):
            self._generic = stft._generic
            # already a Spectrogram (of any stage): the reference keeps the given data and only re-types it
            # (madmom/audio/spectrogram.py:125-127), so a filtered / logarithmic / difference matrix stays what it is
            DeviceResult.__init__(self, stft._chain.copy())
            self._source = stft
            self._output = stft._output
            self.stft = stft.stft
            self.frames = stft.stft.frames
            return
        if not isinstance(stft, ShortTimeFourierTransform):
            stft = ShortTimeFourierTransform(stft, **kwargs)
        DeviceResult.__init__(self, stft._chain.copy())
        self.stft = stft
        self.frames = stft.frames

    def _compute(self):
        if getattr(self, "_source", None) is not None:
            return self._source.tensor
        return DeviceResult._compute(self)

    def _shape(self):
        if getattr(self, "_source", None) is not None:
            return self._source.shape
        return self.stft.num_frames, self.stft.num_bins

    @property
    def num_frames(self):
        return self.shape[0]

    @property
    def num_bins(self):
        return self.shape[1]

    @property
    def bin_frequencies(self):
        return self.stft.bin_frequencies

    def diff(self, **kwargs):
        return SpectrogramDifference(self, **kwargs)

    def filter(self, **kwargs):
        return FilteredSpectrogram(self, **kwargs)

    def log(self, **kwargs):
        return LogarithmicSpectrogram(self, **kwargs)


class SpectrogramProcessor(Processor):
    def __init__(self, **kwargs):
        pass

    def process(self, data, **kwargs):
        return Spectrogram(data, **kwargs)


class FilteredSpectrogram(Spectrogram):
    """Spectrogram times filterbank, (num_frames, num_bands) float32 (banded FP32 contraction in the kernel)."""
    _output = "filt"

    def __init__(self, spectrogram, filterbank=FILTERBANK, num_bands=NUM_BANDS, fmin=FMIN, fmax=FMAX, fref=FREF,
                 norm_filters=NORM_FILTERS, unique_filters=UNIQUE_FILTERS, **kwargs):
        if not isinstance(spectrogram, Spectrogram):
            spectrogram = Spectrogram(spectrogram, **kwargs)
        c0 = spectrogram._chain
        generic = (spectrogram._generic or c0.filterbank is not None or c0.log_mode != 0 or c0.diff_frames != 0 or
                   isinstance(spectrogram, (FilteredSpectrogram, LogarithmicSpectrogram, SpectrogramDifference)))
        if isinstance(filterbank, type) and issubclass(filterbank, Filterbank):
            fb_kw = dict(num_bands=num_bands, fmin=fmin, fmax=fmax, norm_filters=norm_filters,
                         unique_filters=unique_filters)
            if issubclass(filterbank, LogarithmicFilterbank):
                fb_kw["fref"] = fref
            filterbank = filterbank(spectrogram.bin_frequencies, **fb_kw)
        if not isinstance(filterbank, Filterbank):
            raise TypeError("not a Filterbank type or instance: %s" % filterbank)
        if filterbank.shape[0] != spectrogram.num_bins:
            raise ValueError("filterbank has %d bins, the spectrogram %d" % (filterbank.shape[0],
                                                                           spectrogram.num_bins))
        DeviceResult.__init__(self, spectrogram._chain.copy())
        self._generic = generic
        self._filter_source = spectrogram if generic else None
        if not generic:
            self._chain.filterbank = filterbank
        self.filterbank = filterbank
        self.stft = spectrogram.stft
        self.frames = spectrogram.frames

    def _compute(self):
        if self._generic:
            # np.dot(spectrogram, filterbank) (madmom/audio/spectrogram.py:382) on the device, float32 like there
            import torch
            x = self._filter_source.tensor
            w = torch.from_numpy(np.ascontiguousarray(np.asarray(self.filterbank), dtype=np.float32)).to(x.device)
            return torch.matmul(x, w)
        return Spectrogram._compute(self)

    def _shape(self):
        rows = self._filter_source.shape[0] if self._generic else self.stft.num_frames
        return rows, self.filterbank.shape[1]

    @property
    def bin_frequencies(self):
        return self.filterbank.center_frequencies


class FilteredSpectrogramProcessor(Processor):
    """madmom/audio/spectrogram.py:406-470; caches the filterbank after the first call."""

    def __init__(self, filterbank=FILTERBANK, num_bands=NUM_BANDS, fmin=FMIN, fmax=FMAX, fref=FREF,
                 norm_filters=NORM_FILTERS, unique_filters=UNIQUE_FILTERS, **kwargs):
        self.filterbank = filterbank
        self.num_bands = num_bands
        self.fmin = fmin
        self.fmax = fmax
        self.fref = fref
        self.norm_filters = norm_filters
        self.unique_filters = unique_filters

    def process(self, data, **kwargs):
        args = dict(filterbank=self.filterbank, num_bands=self.num_bands, fmin=self.fmin, fmax=self.fmax,
                    fref=self.fref, norm_filters=self.norm_filters, unique_filters=self.unique_filters)
        args.update(kwargs)
        data = FilteredSpectrogram(data, **args)
        self.filterbank = data.filterbank
        return data


class LogarithmicSpectrogram(Spectrogram):
    """log(mul * x + add) of a (filtered) spectrogram, float32, evaluated in that order like the reference
    (madmom/audio/spectrogram.py:519-544)."""
    _output = "log"

    def __init__(self, spectrogram, log=LOG, mul=MUL, add=ADD, **kwargs):
        if not isinstance(spectrogram, Spectrogram):
            spectrogram = Spectrogram(spectrogram, **kwargs)
        generic = (spectrogram._generic or spectrogram._chain.log_mode != 0 or spectrogram._chain.diff_frames != 0 or
                   isinstance(spectrogram, (LogarithmicSpectrogram, SpectrogramDifference)))
        DeviceResult.__init__(self, spectrogram._chain.copy())
        self._generic = generic
        self._log = log
        c = self._chain
        mode = _engine.log_mode_of(log, mul, add)            # raises for a logarithm the kernels do not have
        if not generic:
            c.log_mode = mode
            c.mul = 1.0 if mul is None else float(mul)
            c.add = 0.0 if add is None else float(add)
        # an unfiltered spectrogram is scaled in the magnitude epilogue of the kernel ('spec' output)
        self._output = "log" if c.filterbank is not None else "spec"
        self.mul = mul
        self.add = add
        self.stft = spectrogram.stft
        self.frames = spectrogram.frames
        self.spectrogram = spectrogram

    def _compute(self):
        if self._generic:
            # log(mul * x + add) in float32, in that order (madmom/audio/spectrogram.py:537-544)
            import torch
            x = self.spectrogram.tensor
            if self.mul is not None:
                x = x * float(self.mul)
            if self.add is not None:
                x = x + float(self.add)
            if self._log is None:
                return x
            return {np.log10: torch.log10, np.log: torch.log, np.log2: torch.log2}[self._log](x)
        return Spectrogram._compute(self)

    def _shape(self):
        return self.spectrogram.shape

    @property
    def filterbank(self):
        return self.spectrogram.filterbank

    @property
    def bin_frequencies(self):
        return self.spectrogram.bin_frequencies


class LogarithmicSpectrogramProcessor(Processor):
    def __init__(self, log=LOG, mul=MUL, add=ADD, **kwargs):
        self.log = log
        self.mul = mul
        self.add = add

    def process(self, data, **kwargs):
        args = dict(log=self.log, mul=self.mul, add=self.add)
        args.update(kwargs)
        return LogarithmicSpectrogram(data, **args)


class LogarithmicFilteredSpectrogram(LogarithmicSpectrogram, FilteredSpectrogram):
    """FilteredSpectrogram followed by LogarithmicSpectrogram (madmom/audio/spectrogram.py:664-748)."""

    def __init__(self, spectrogram, **kwargs):
        mul = kwargs.pop("mul", MUL)
        add = kwargs.pop("add", ADD)
        log = kwargs.pop("log", LOG)
        if not isinstance(spectrogram, FilteredSpectrogram):
            spectrogram = FilteredSpectrogram(spectrogram, **kwargs)
        LogarithmicSpectrogram.__init__(self, spectrogram, log=log, mul=mul, add=add)

    @property
    def filterbank(self):
        return self.spectrogram.filterbank

    @property
    def bin_frequencies(self):
        return self.filterbank.center_frequencies


class LogarithmicFilteredSpectrogramProcessor(Processor):
    """madmom/audio/spectrogram.py:751-824"""

    def __init__(self, filterbank=FILTERBANK, num_bands=NUM_BANDS, fmin=FMIN, fmax=FMAX, fref=FREF,
                 norm_filters=NORM_FILTERS, unique_filters=UNIQUE_FILTERS, mul=MUL, add=ADD, **kwargs):
        self.filterbank = filterbank
        self.num_bands = num_bands
        self.fmin = fmin
        self.fmax = fmax
        self.fref = fref
        self.norm_filters = norm_filters
        self.unique_filters = unique_filters
        self.mul = mul
        self.add = add

    def process(self, data, **kwargs):
        args = dict(filterbank=self.filterbank, num_bands=self.num_bands, fmin=self.fmin, fmax=self.fmax,
                    fref=self.fref, norm_filters=self.norm_filters, unique_filters=self.unique_filters,
                    mul=self.mul, add=self.add)
        args.update(kwargs)
        data = LogarithmicFilteredSpectrogram(data, **args)
        self.filterbank = data.filterbank
        return data


class SpectrogramDifference(Spectrogram):
    """Lagged (optionally maximum-filtered, optionally half-wave rectified) difference of a filtered /
    logarithmic filtered spectrogram; `keep_dims=True` keeps the first `diff_frames` rows as zeros
    (madmom/audio/spectrogram.py:960-1004)."""
    _output = "diff"

    def __init__(self, spectrogram, diff_ratio=DIFF_RATIO, diff_frames=DIFF_FRAMES, diff_max_bins=DIFF_MAX_BINS,
                 positive_diffs=POSITIVE_DIFFS, keep_dims=True, **kwargs):
        if not isinstance(spectrogram, Spectrogram):
            spectrogram = Spectrogram(spectrogram, **kwargs)
        generic = (spectrogram._generic or isinstance(spectrogram, SpectrogramDifference) or
                   spectrogram._chain.diff_frames != 0)
        if diff_frames is None:
            diff_frames = _diff_frames(diff_ratio, hop_size=spectrogram.stft.frames.hop_size,
                                       frame_size=spectrogram.stft.frames.frame_size,
                                       window=spectrogram.stft.window)
        diff_frames = int(diff_frames)
        if diff_frames < 1:
            raise ValueError("number of `diff_frames` must be >= 1")
        DeviceResult.__init__(self, spectrogram._chain.copy())
        self._generic = generic
        c = self._chain
        c.diff_frames = diff_frames
        c.diff_max_bins = int(diff_max_bins) if diff_max_bins is not None and diff_max_bins > 1 else 0
        c.positive_diffs = bool(positive_diffs)
        self.spectrogram = spectrogram
        self.stft = spectrogram.stft
        self.frames = spectrogram.frames
        self.diff_ratio = diff_ratio
        self.diff_frames = diff_frames
        self.diff_max_bins = diff_max_bins
        self.positive_diffs = positive_diffs
        self.keep_dims = bool(keep_dims)

    @property
    def _fused(self):
        # unfiltered spectrograms (1024+ columns) do not fit the fused kernel's shared-memory ring, and generic
        # sources are not described by the chain: both are materialised and differenced by mmb_matrix_diff
        return self._chain.filterbank is not None and not self._generic

    def _run(self, want_diff, want_odf):
        if self._fused:
            names = (("diff",) if want_diff else ()) + (("odf",) if want_odf else ())
            out = self._chain.run(names)
            return out.get("diff"), out.get("odf")
        from .. import _lib
        c = self._chain
        return _lib.matrix_diff(self.spectrogram.tensor, c.diff_frames, c.diff_max_bins, c.positive_diffs,
                                want_diff=want_diff, want_odf=want_odf)

    def _compute(self):
        t = self._run(True, False)[0]
        return t if self.keep_dims else t[self.diff_frames:]

    def _shape(self):
        rows, cols = self.spectrogram.shape
        return (rows if self.keep_dims else max(rows - self.diff_frames, 0)), cols

    @property
    def filterbank(self):
        return self.spectrogram.filterbank

    @property
    def bin_frequencies(self):
        return self.spectrogram.bin_frequencies

    def positive_diff(self):
        return np.maximum(self.numpy(), 0)

    def flux(self):
        """Row sums of the difference as a device vector (the `spectral_flux` / `superflux` reduction,
        madmom/features/onsets.py:163-240), fused into the same kernel pass."""
        return _RowSum(self)


class _RowSum(DeviceResult):
    _output = "odf"

    def __init__(self, diff):
        DeviceResult.__init__(self, diff._chain.copy())
        self.diff = diff

    def _compute(self):
        t = self.diff._run(False, True)[1]
        return t if self.diff.keep_dims else t[self.diff.diff_frames:]

    def _shape(self):
        return (self.diff.shape[0],)


class _StackedDifference(DeviceResult):
    """np.hstack((spectrogram, difference)) written in place by the kernel: (num_frames, 2 * num_bands)."""
    _output = "stack"

    def __init__(self, diff):
        DeviceResult.__init__(self, diff._chain.copy())
        self.diff = diff
        self.spectrogram = diff.spectrogram
        self.stft = diff.stft
        self.frames = diff.frames
        self.diff_frames = diff.diff_frames

    def _compute(self):
        if self.diff._fused:
            return DeviceResult._compute(self)
        import torch
        return torch.cat((self.spectrogram.tensor, self.diff.tensor), dim=1)

    def _shape(self):
        rows, cols = self.spectrogram.shape
        return rows, 2 * cols

    @property
    def num_frames(self):
        return self.shape[0]

    @property
    def num_bins(self):
        return self.shape[1]


class SpectrogramDifferenceProcessor(Processor):
    """madmom/audio/spectrogram.py:1025-1144.  `process(data, reset=True)` returns the difference with the
    first `diff_frames` rows zero (the reference gets there by prepending +inf rows, :1120-1133), optionally
    stacked with the spectrogram.  Continuing a stream (`reset=False`) is not available on the GPU path."""

    def __init__(self, diff_ratio=DIFF_RATIO, diff_frames=DIFF_FRAMES, diff_max_bins=DIFF_MAX_BINS,
                 positive_diffs=POSITIVE_DIFFS, stack_diffs=None, **kwargs):
        self.diff_ratio = diff_ratio
        self.diff_frames = diff_frames
        self.diff_max_bins = diff_max_bins
        self.positive_diffs = positive_diffs
        self.stack_diffs = stack_diffs
        self._buffer = None

    def __getstate__(self):
        state = self.__dict__.copy()
        state.pop("_buffer", None)
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)
        self._buffer = None

    def process(self, data, reset=True, **kwargs):
        if not reset:
            raise NotImplementedError("SpectrogramDifferenceProcessor(reset=False) (frame-by-frame streaming) is "
                                      "not available on the GPU path")
        args = dict(diff_ratio=self.diff_ratio, diff_frames=self.diff_frames, diff_max_bins=self.diff_max_bins,
                    positive_diffs=self.positive_diffs)
        args.update(kwargs)
        args.pop("keep_dims", None)
        diff = SpectrogramDifference(data, keep_dims=True, **args)
        if self.stack_diffs is None:
            return diff
        if self.stack_diffs is np.hstack:
            return _StackedDifference(diff)
        # any other stacking function only moves data: apply it to the two device results' host copies
        return self.stack_diffs((np.asarray(diff.spectrogram), np.asarray(diff)))

    def reset(self):
        self._buffer = None


class SuperFluxProcessor(SequentialProcessor):
    """STFT -> 24 bands/octave un-normalised log filterbank -> log10(x + 1) -> maximum filter 3 + positive
    difference at ratio 0.5 (madmom/audio/spectrogram.py:1230-1262).  One fused kernel launch."""

    def __init__(self, **kwargs):
        stft = ShortTimeFourierTransformProcessor(**kwargs)
        spec = SpectrogramProcessor(**kwargs)
        filt = FilteredSpectrogramProcessor(num_bands=24, norm_filters=False, **kwargs)
        log = LogarithmicSpectrogramProcessor(**kwargs)
        diff = SpectrogramDifferenceProcessor(diff_ratio=0.5, diff_max_bins=3, positive_diffs=True, **kwargs)
        super(SuperFluxProcessor, self).__init__([stft, spec, filt, log, diff])


class MultiFrameProcessor(Processor):
    """Multi-resolution stack in one device matrix: for every frame size the chain
    FramedSignal -> STFT -> filterbank -> log -> hstack((spec, positive diff)), all blocks side by side.

    New convenience class (the reference spells this ParallelProcessor([...]) + np.hstack,
    madmom/features/onsets.py:778-791, madmom/features/beats.py:93-106; that spelling works too).  The kernels
    write straight into the column blocks of the (num_frames, sum 2 B_i) result, the PCM is uploaded once."""

    def __init__(self, frame_sizes=(1024, 2048, 4096), fps=100, num_bands=6, fmin=30., fmax=17000.,
                 norm_filters=True, mul=1., add=1., diff_ratio=0.5, positive_diffs=True, **kwargs):
        self.frame_sizes = tuple(frame_sizes)
        self.fps = fps
        self.num_bands = num_bands
        self.fmin, self.fmax = fmin, fmax
        self.norm_filters = norm_filters
        self.mul, self.add = mul, add
        self.diff_ratio = diff_ratio
        self.positive_diffs = positive_diffs

    def process(self, data, **kwargs):
        import torch
        from .signal import FramedSignal, Signal
        if not isinstance(data, Signal):
            data = Signal(data, **kwargs)
        stacks, shared = [], None
        for n in self.frame_sizes:
            frames = FramedSignal(data, frame_size=n, fps=self.fps)
            if shared is not None:
                frames._device_pcm = shared          # one upload for all resolutions
            spec = LogarithmicFilteredSpectrogram(frames, num_bands=self.num_bands, fmin=self.fmin,
                                                  fmax=self.fmax, norm_filters=self.norm_filters, mul=self.mul,
                                                  add=self.add)
            diff = SpectrogramDifference(spec, diff_ratio=self.diff_ratio, positive_diffs=self.positive_diffs)
            stacks.append(_StackedDifference(diff))
            shared = frames.device_pcm(_engine.current_device())
        out = _Concatenated(stacks)
        return out


class _Concatenated(DeviceResult):
    """Column-wise concatenation of device results (kept on the GPU)."""

    def __init__(self, parts):
        DeviceResult.__init__(self, None)
        self.parts = parts

    def _compute(self):
        import torch
        return torch.cat([p.tensor for p in self.parts], dim=1)

    def _shape(self):
        return self.parts[0].shape[0], sum(p.shape[1] for p in self.parts)
