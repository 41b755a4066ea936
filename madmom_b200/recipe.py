"""Host-side scalar logic of the front end and the named parameter sets ("recipes") of the reference.

Everything here is integer / float64 metadata that the reference also computes once on the host; it must be
bit-exact, so it is done in Python floats exactly as the reference does and passed to the CUDA library as
integers or float64 (never recomputed in FP32 on the device).
"""
import numpy as np

from .audio.filters import LogarithmicFilterbank, Filterbank

SAMPLE_RATE = 44100


def resolve_origin(origin, frame_size):
    """Literal origin -> int (FramedSignal.__init__, madmom/audio/signal.py:1126-1142)."""
    if origin in ("center", "offline"):
        origin = 0
    elif origin in ("left", "past", "online"):
        origin = (frame_size - 1) / 2
    elif origin in ("right", "future", "stream"):
        origin = -(frame_size / 2)
    return int(origin)


def num_frames(num_samples, hop_size, end="normal"):
    """ceil(len / hop) ('normal') or floor(len / hop + 1) ('extend') in float64
    (madmom/audio/signal.py:1145-1156)."""
    if end == "extend":
        return int(np.floor(num_samples / float(hop_size) + 1))
    if end == "normal":
        return int(np.ceil(num_samples / float(hop_size)))
    raise ValueError("end of signal handling '%s' unknown" % end)


def frame_start(index, frame_size, hop_size, origin=0):
    """int(index * hop) - frame_size // 2 - origin (signal_frame, madmom/audio/signal.py:917-919)."""
    return int(index * hop_size) - int(frame_size) // 2 - int(origin)


_NUMPY_WINDOWS = (np.hanning, np.hamming, np.blackman, np.bartlett)
_window_cache = {}


def window_of(window, frame_size):
    """`window(frame_size)` for a callable window.  numpy's own window functions are pure, so their (read-only)
    results are kept: a per-file call of the drop-in processors does not rebuild the window every time."""
    if not callable(window):
        return window
    if not any(window is w for w in _NUMPY_WINDOWS):
        return window(frame_size)
    key = (window.__name__, int(frame_size))
    w = _window_cache.get(key)
    if w is None:
        w = window(frame_size)
        w.setflags(write=False)
        _window_cache[key] = w
    return w


def fft_window(window, frame_size, signal_dtype):
    """float64 window actually multiplied onto the samples: integer PCM is left un-scaled and the window
    divided by iinfo(dtype).max instead (ShortTimeFourierTransform.__new__, madmom/audio/stft.py:331-349).
    window=None on float PCM means a rectangular window (the kernel always multiplies)."""
    window = window_of(window, frame_size)
    if window is not None:
        window = np.asarray(window, dtype=np.float64)
        if window.shape != (frame_size,):
            raise ValueError("window must have frame_size (%d) elements" % frame_size)
    if np.issubdtype(np.dtype(signal_dtype), np.integer):
        scale = float(np.iinfo(signal_dtype).max)
        return (np.ones(frame_size) if window is None else window) / scale
    return np.ones(frame_size) if window is None else window


def fft_frequencies(num_fft_bins, sample_rate):
    """madmom/audio/stft.py:29-46"""
    return np.fft.fftfreq(num_fft_bins * 2, 1. / sample_rate)[:num_fft_bins]


def diff_frames(diff_ratio, hop_size, frame_size, window=np.hanning):
    """_diff_frames, madmom/audio/spectrogram.py:834-864: first window sample above ratio * max, distance
    to the window centre in hops, Python (banker's) round, at least 1."""
    window = window_of(window, frame_size)
    # the reference takes Python's max() of the array; np.max returns the same float64 without the per-element loop
    sample = np.argmax(window > float(diff_ratio) * np.max(window))
    diff_samples = len(window) / 2 - sample
    return int(max(1, round(diff_samples / hop_size)))


class Recipe(object):
    """One front-end parameter set (SURVEY.md section 8a recipe table).  `diff_frames` None -> from
    `diff_ratio`; `diff_max_bins` None/1 -> no frequency maximum filter."""

    def __init__(self, frame_size=2048, hop_size=441., fps=None, origin=0, end="normal", window=np.hanning,
                 circular_shift=False, filterbank=LogarithmicFilterbank, num_bands=12, fmin=30., fmax=17000.,
                 fref=440., norm_filters=True, unique_filters=True, log=np.log10, mul=1., add=1.,
                 diff_ratio=0.5, diff_frames=None, diff_max_bins=None, positive_diffs=True,
                 sample_rate=SAMPLE_RATE):
        if fps is not None:
            hop_size = sample_rate / float(fps)                 # madmom/audio/signal.py:1121-1123
        self.frame_size, self.hop_size = int(frame_size), float(hop_size)
        self.origin, self.end, self.window, self.circular_shift = origin, end, window, bool(circular_shift)
        self.filterbank, self.num_bands, self.fmin, self.fmax, self.fref = filterbank, num_bands, fmin, fmax, fref
        self.norm_filters, self.unique_filters = norm_filters, unique_filters
        self.log, self.mul, self.add = log, mul, add
        self.diff_ratio, self.diff_frames = diff_ratio, diff_frames
        self.diff_max_bins, self.positive_diffs = diff_max_bins, positive_diffs
        self.sample_rate = sample_rate

    @property
    def fps(self):
        return self.sample_rate / self.hop_size

    def replace(self, **kw):
        d = dict(self.__dict__)
        d.update(kw)
        return Recipe(**d)

    def window_array(self):
        w = self.window(self.frame_size) if callable(self.window) else self.window
        return np.ones(self.frame_size) if w is None else np.asarray(w, dtype=np.float64)

    def build_filterbank(self):
        fb = self.filterbank
        if fb is None or isinstance(fb, Filterbank):
            return fb
        if isinstance(fb, type) and issubclass(fb, Filterbank):
            binf = fft_frequencies(self.frame_size // 2, self.sample_rate)
            kw = dict(num_bands=self.num_bands, fmin=self.fmin, fmax=self.fmax, norm_filters=self.norm_filters,
                      unique_filters=self.unique_filters)
            if issubclass(fb, LogarithmicFilterbank):
                kw["fref"] = self.fref
            return fb(binf, **kw)
        raise TypeError("not a Filterbank type or instance: %s" % fb)

    def resolved_diff_frames(self):
        if self.diff_frames is not None:
            return int(self.diff_frames)
        return diff_frames(self.diff_ratio, self.hop_size, self.frame_size, self.window_array())


# bin/SuperFlux:59-69 uses fps 200; SuperFluxProcessor + default FramedSignal is fps 100 (config 1)
SUPERFLUX = Recipe(frame_size=2048, hop_size=441., num_bands=24, norm_filters=False, diff_ratio=0.5,
                   diff_max_bins=3, positive_diffs=True)
SUPERFLUX_200 = SUPERFLUX.replace(hop_size=220.5)
# bin/LogFiltSpecFlux:60-69 (config 2)
LOGFILTSPECFLUX = Recipe(frame_size=2048, hop_size=441., num_bands=12, norm_filters=False, diff_ratio=0.5,
                         diff_max_bins=None, positive_diffs=True)
# bin/ComplexFlux:60-74 (config 5)
COMPLEXFLUX = SUPERFLUX_200.replace(circular_shift=True)


def multires(mul, diff_ratio, fps=100, frame_sizes=(1024, 2048, 4096)):
    """RNNOnsetProcessor (mul 5, ratio .25; madmom/features/onsets.py:767-791) / RNNBeatProcessor
    (mul 1, ratio .5; madmom/features/beats.py:80-106) pre-processing: one recipe per frame size."""
    return [Recipe(frame_size=n, fps=fps, num_bands=6, norm_filters=True, mul=mul, add=1, diff_ratio=diff_ratio,
                   positive_diffs=True) for n in frame_sizes]


RNN_ONSET = multires(5, 0.25)
RNN_BEAT = multires(1, 0.5)
