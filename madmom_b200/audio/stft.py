"""Short-time Fourier transform stage.

Mirror of madmom/audio/stft.py: `fft_frequencies` (:29-46), `ShortTimeFourierTransform` (:202-426) and
`ShortTimeFourierTransformProcessor` (:432-549).

Unlike the reference the transform is *lazy*: constructing the object only records the chain (frames, window,
circular shift); the complex bins are computed by the fused CUDA kernel when somebody actually needs them
(`.tensor`, `np.asarray(...)`, indexing).  A chain that continues to a spectrogram / filterbank / log / difference
never materialises the (T, K) complex matrix at all.
"""
import numpy as np

from ..processors import Processor
from ..recipe import fft_frequencies, window_of  # noqa: F401  (fft_frequencies re-exported, same name as the reference)
from .._engine import Chain
from .signal import FramedSignal

STFT_DTYPE = np.complex64


class DeviceResult(object):
    """A matrix that lives on the GPU and is computed on first use.  Behaves like a read-only ndarray for the
    consumers the reference has (np.asarray, np.hstack, indexing, .shape, arithmetic through numpy)."""
    _output = None          # name of the kernel output this object represents

    def __init__(self, chain):
        self._chain = chain
        self._tensor = None
        self._host = None

    # ---- materialisation
    def _compute(self):
        return self._chain.run((self._output,))[self._output]

    @property
    def tensor(self):
        """CUDA tensor (computed on first access)."""
        if self._tensor is None:
            self._tensor = self._compute()
        return self._tensor

    def numpy(self):
        if self._host is None:
            self._host = self.tensor.cpu().numpy()
        return self._host

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        if dtype is not None and np.dtype(dtype) != a.dtype:
            return a.astype(dtype)
        return a.copy() if copy else a

    def __getstate__(self):
        state = dict(self.__dict__)
        state["_tensor"] = None      # CUDA handles do not travel through pickle; the host copy does
        if self._tensor is not None and self._host is None:
            state["_host"] = self.numpy()
        return state

    # ---- ndarray look-alike
    def _shape(self):
        raise NotImplementedError

    @property
    def shape(self):
        return self._shape()

    @property
    def ndim(self):
        return len(self.shape)

    @property
    def size(self):
        return int(np.prod(self.shape))

    @property
    def dtype(self):
        return np.dtype(np.float32)

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, index):
        return self.numpy()[index]

    def __iter__(self):
        return iter(self.numpy())

    def __getattr__(self, name):
        # ndarray methods (sum, max, T, astype, ...) are answered by the host copy
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self.numpy(), name)

    def __repr__(self):
        return "<%s %s on %s>" % (type(self).__name__, "x".join(str(s) for s in self.shape),
                                  "GPU" if self._tensor is not None else "GPU (lazy)")


def _binary(op):
    def f(self, other):
        return getattr(self.numpy(), op)(np.asarray(other))
    return f


for _op in ("__add__", "__sub__", "__mul__", "__truediv__", "__radd__", "__rsub__", "__rmul__", "__rtruediv__",
            "__lt__", "__le__", "__gt__", "__ge__", "__eq__", "__ne__", "__pow__"):
    setattr(DeviceResult, _op, _binary(_op))
DeviceResult.__hash__ = object.__hash__


class ShortTimeFourierTransform(DeviceResult):
    """Complex STFT of a FramedSignal, (num_frames, frame_size / 2) complex64."""
    _output = "stft"

    def __init__(self, frames, window=np.hanning, fft_size=None, circular_shift=False, include_nyquist=False,
                 fft_window=None, **kwargs):
        if isinstance(frames, ShortTimeFourierTransform):
            frames = frames.frames
        if not isinstance(frames, FramedSignal):
            frames = FramedSignal(frames, **kwargs)
        if frames.ndim != 2:
            # the reference raises the same for multi-channel frames (madmom/audio/stft.py:81-84)
            raise ValueError("frames must be a 2D array or iterable, got %s with shape %s." %
                             (type(frames), frames.shape))
        frame_size = frames.shape[1]
        if fft_size is not None and fft_size != frame_size:
            raise NotImplementedError("fft_size != frame_size (zero padding) is not available on the GPU path")
        if frame_size not in (1024, 2048, 4096):
            raise NotImplementedError("frame_size must be 1024, 2048 or 4096 on the GPU path, not %d" % frame_size)
        window = window_of(window, frame_size)
        chain = Chain(frames, window=window, circular_shift=circular_shift, fft_window=fft_window,
                      include_nyquist=include_nyquist)
        DeviceResult.__init__(self, chain)
        self.frames = frames
        self.window = window
        self.fft_window = chain.resolved_fft_window()
        if window is None and fft_window is None and not np.issubdtype(frames.signal.dtype, np.integer):
            # stft.py:331-350: no window and no integer range to scale by -> the attribute is None (the kernel still
            # multiplies by the ones the chain resolved)
            self.fft_window = None
        self.fft_size = frame_size
        self.circular_shift = bool(circular_shift)
        self.include_nyquist = bool(include_nyquist)

    def _shape(self):
        return self.frames.num_frames, self.num_bins

    @property
    def dtype(self):
        return np.dtype(STFT_DTYPE)

    @property
    def num_frames(self):
        return self.frames.num_frames

    @property
    def num_bins(self):
        # stft.py:93-95: fft_size >> 1, plus the Nyquist bin on request
        return (self.fft_size >> 1) + (1 if self.include_nyquist else 0)

    @property
    def bin_frequencies(self):
        # stft.py:386-389: fft_frequencies(num_bins, sample_rate) with num_bins as stored (Nyquist included)
        return fft_frequencies(self.num_bins, self.frames.signal.sample_rate)

    def spec(self, **kwargs):
        from .spectrogram import Spectrogram
        return Spectrogram(self, **kwargs)

    def phase(self, **kwargs):
        return Phase(self, **kwargs)


STFT = ShortTimeFourierTransform


def phase(stft):
    """Phase of a complex STFT, np.angle (madmom/audio/stft.py:136-151).  Device results stay on the device."""
    if isinstance(stft, ShortTimeFourierTransform):
        return Phase(stft)
    return np.angle(stft)


def local_group_delay(phase):
    """Local group delay of a phase matrix: unwrap along the bins, lgd[:, k] = up[k] - up[k + 1], last bin 0
    (madmom/audio/stft.py:154-179).  A `Phase` object is answered on the device, a plain array on the host."""
    if isinstance(phase, Phase):
        return LocalGroupDelay(phase)
    phase = np.asarray(phase)
    if phase.ndim != 2:
        raise ValueError("phase must be a 2D array")
    up = np.unwrap(phase)
    up[:, :-1] -= up[:, 1:]
    up[:, -1] = 0
    return up


lgd = local_group_delay


class Phase(DeviceResult):
    """np.angle of the STFT, (num_frames, num_bins) float32 (madmom/audio/stft.py:556-642).  Built from anything a
    ShortTimeFourierTransform can be built from; `circular_shift` then defaults to True, and an STFT without it
    draws the reference's RuntimeWarning."""

    def __init__(self, stft, **kwargs):
        import warnings
        if isinstance(stft, Phase):
            stft = stft.stft
        if not isinstance(stft, ShortTimeFourierTransform):
            circular_shift = kwargs.pop("circular_shift", True)
            stft = ShortTimeFourierTransform(stft, circular_shift=circular_shift, **kwargs)
        if not stft.circular_shift:
            warnings.warn("`circular_shift` of the STFT must be set to 'True' for correct phase", RuntimeWarning)
        DeviceResult.__init__(self, stft._chain.copy())
        self.stft = stft

    def _compute(self):
        from .. import _lib
        return _lib.phase_lgd(self.stft.tensor, want_phase=True, want_lgd=False)[0]

    def _shape(self):
        return self.stft.shape

    @property
    def bin_frequencies(self):
        return self.stft.bin_frequencies

    @property
    def num_frames(self):
        return int(self.shape[0])

    @property
    def num_bins(self):
        return int(self.shape[1])

    def local_group_delay(self, **kwargs):
        return LocalGroupDelay(self, **kwargs)

    lgd = local_group_delay


class LocalGroupDelay(DeviceResult):
    """Local group delay of the phase of an STFT (madmom/audio/stft.py:646-708): unwrapped phase differences along
    the bins, last bin 0.  Computed from the complex bins directly (the difference of neighbouring unwrapped phases
    is the wrapped phase difference), so it does not carry the rounding of the reference's float32 cumulative sum
    (up to about 4e-5)."""

    def __init__(self, phase, **kwargs):
        import warnings
        if not isinstance(phase, Phase):
            phase = Phase(phase, circular_shift=True, **kwargs)
        if not phase.stft.circular_shift:
            warnings.warn("`circular_shift` of the STFT must be set to 'True' for correct local group delay")
        DeviceResult.__init__(self, phase._chain.copy())
        self.phase = phase
        self.stft = phase.stft

    def _compute(self):
        from .. import _lib
        return _lib.phase_lgd(self.stft.tensor, want_phase=False, want_lgd=True)[1]

    def _shape(self):
        return self.stft.shape

    @property
    def bin_frequencies(self):
        return self.stft.bin_frequencies

    @property
    def num_frames(self):
        return int(self.shape[0])

    @property
    def num_bins(self):
        return int(self.shape[1])


LGD = LocalGroupDelay


class ShortTimeFourierTransformProcessor(Processor):
    """madmom/audio/stft.py:432-549; the scaled window is cached after the first call like the reference."""

    def __init__(self, window=np.hanning, fft_size=None, circular_shift=False, include_nyquist=False, **kwargs):
        self.window = window
        self.fft_size = fft_size
        self.circular_shift = circular_shift
        self.include_nyquist = include_nyquist
        self.fft_window = None

    def process(self, data, **kwargs):
        data = ShortTimeFourierTransform(data, window=self.window, fft_size=self.fft_size,
                                         circular_shift=self.circular_shift,
                                         include_nyquist=self.include_nyquist, fft_window=self.fft_window, **kwargs)
        self.fft_window = data.fft_window
        return data


STFTProcessor = ShortTimeFourierTransformProcessor
