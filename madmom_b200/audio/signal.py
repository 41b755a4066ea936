"""Signal and (lazy) framing.

Mirror of madmom/audio/signal.py for the part the hot path touches: `Signal` (:490-732), `signal_frame`
(:860-962), `FramedSignal` (:965-1260) and `FramedSignalProcessor` (:1263-1358).

Framing is never materialised here: `FramedSignal` is the descriptor (frame size, hop, integer origin, frame
count) that the CUDA library consumes; frame `i` starts at sample `int(i * hop_size) - frame_size // 2 - origin`
and samples outside the signal read as zero.  All of that integer arithmetic is done in Python floats exactly as
the reference does.  Indexing a single frame on the host (`frames[i]`) is plain slicing and kept for API
compatibility.
"""
import numpy as np

from ..processors import Processor
from ..recipe import resolve_origin, num_frames as _num_frames, frame_start

SAMPLE_RATE = None
NUM_CHANNELS = None
FRAME_SIZE = 2048
HOP_SIZE = 441.
FPS = None
ORIGIN = 0
END_OF_SIGNAL = "normal"
NUM_FRAMES = None


def load_wave_file(filename, sample_rate=None, num_channels=None, start=None, stop=None, dtype=None):
    """Read a RIFF/WAVE file with scipy (madmom/io/audio.py:594-667).  Returns (ndarray, sample_rate).
    Re-sampling is not part of the front end: a `sample_rate` different from the file's raises."""
    from scipy.io import wavfile
    file_rate, data = wavfile.read(filename, mmap=True)
    if sample_rate is not None and sample_rate != file_rate:
        raise NotImplementedError("re-sampling (%s -> %s Hz) is outside the GPU front end" % (file_rate, sample_rate))
    if start is not None or stop is not None:
        a = None if start is None else int(start * file_rate)
        b = None if stop is None else min(len(data), int(stop * file_rate))
        data = data[a:b]
    if dtype is not None and np.dtype(dtype) != data.dtype:
        data = rescale(data, dtype)
    if num_channels is not None:
        data = remix(data, num_channels)
    return data, file_rate


def rescale(signal, dtype=np.float32):
    """Integer PCM -> float in [-1, 1] by dividing by iinfo.max (madmom/audio/signal.py:266-292)."""
    if not np.issubdtype(np.dtype(dtype), np.floating):
        raise ValueError("only float dtypes are supported, not %s." % dtype)
    if np.issubdtype(signal.dtype, np.floating):
        return signal.astype(dtype)
    if not np.issubdtype(signal.dtype, np.integer):
        raise ValueError("only integer and float signals can be rescaled, not %s." % signal.dtype)
    return signal.astype(dtype) / np.iinfo(signal.dtype).max


def _db_to_factor(gain):
    """dB -> amplitude factor as the reference spells it: sqrt(10) ** (gain / 10) (madmom/audio/signal.py:96)."""
    return np.power(np.sqrt(10.), 0.1 * gain)


def adjust_gain(signal, gain):
    """Scale by `gain` dB and cast back to the input dtype (integer PCM truncates towards zero); amplifying integer
    PCM is refused because it could clip (madmom/audio/signal.py:71-104)."""
    factor = _db_to_factor(gain)
    if factor > 1 and np.issubdtype(signal.dtype, np.integer):
        raise ValueError("positive gain adjustments are only supported for float dtypes.")
    return np.asanyarray(signal * factor, dtype=signal.dtype)


def attenuate(signal, attenuation):
    """`adjust_gain` by -attenuation dB; 0 dB returns the signal itself (madmom/audio/signal.py:107-131)."""
    return signal if attenuation == 0 else adjust_gain(signal, -attenuation)


def normalize(signal):
    """Scale to the full range of the dtype: floats to [-1, 1], int16 / int32 to +-iinfo.max; same dtype out
    (madmom/audio/signal.py:134-167)."""
    scaling = float(np.max(np.abs(signal)))
    if np.issubdtype(signal.dtype, np.integer):
        if signal.dtype not in (np.int16, np.int32):
            raise ValueError("only float and np.int16/32 dtypes supported, not %s." % signal.dtype)
        scaling /= np.iinfo(signal.dtype).max
    return np.asanyarray(signal / scaling, dtype=signal.dtype)


def smooth(signal, kernel):
    """'same'-sized convolution of a 1-D signal, or of every channel of a 2-D one, with `kernel`: an array, or a
    length > 1 for a Hamming window of that length; None or 0 leaves the signal as it is
    (madmom/audio/signal.py:22-68; `peak_picking` smooths activations this way, on the device inside
    mmb_peak_picking)."""
    if kernel is None:
        return signal
    if isinstance(kernel, (int, np.integer)):
        if kernel == 0:
            return signal
        if kernel < 2:
            raise ValueError("can't create a smoothing kernel of size %d" % kernel)
        kernel = np.hamming(kernel)
    elif not (isinstance(kernel, np.ndarray) and kernel.ndim == 1 and len(kernel) > 1):
        raise ValueError("can't smooth signal with %s" % (kernel,))
    x = np.asarray(signal)
    if x.ndim == 1:
        return np.convolve(x, kernel, "same")
    if x.ndim == 2:
        return np.stack([np.convolve(x[:, c], kernel, "same") for c in range(x.shape[1])], axis=1)
    raise ValueError("signal must be either 1D or 2D")


def trim(signal, where="fb"):
    """Cut leading ('f') and / or trailing ('b') sample frames in which every channel is zero
    (madmom/audio/signal.py:295-329)."""
    where = where.upper()
    x = np.asarray(signal)
    nonzero = np.nonzero(x if x.ndim == 1 else np.abs(x).sum(axis=tuple(range(1, x.ndim))))[0]
    first, last = 0, len(x)
    if len(nonzero):
        if "F" in where:
            first = int(nonzero[0])
        if "B" in where:
            last = int(nonzero[-1]) + 1
    elif "F" in where:
        first = last                                         # all zeros: nothing is left, like np.trim_zeros
    return signal[first:last]


def energy(signal):
    """Sum of the squared samples in float64, all channels together (madmom/audio/signal.py:332-365); one value per
    frame for a FramedSignal."""
    if isinstance(signal, FramedSignal):
        return np.array([energy(frame) for frame in signal])
    x = np.asarray(signal)
    if np.iscomplexobj(x):
        x = np.abs(x)
    x = x.astype(float).ravel()
    return float(np.dot(x, x))


def root_mean_square(signal):
    """sqrt(energy / number of samples) (madmom/audio/signal.py:368-392)."""
    if isinstance(signal, FramedSignal):
        return np.array([root_mean_square(frame) for frame in signal])
    return float(np.sqrt(energy(signal) / np.asarray(signal).size))


def sound_pressure_level(signal, p_ref=None):
    """20 log10(rms / p_ref) in dB; p_ref defaults to the largest value of an integer sample type, 1 for floats;
    silence gives the most negative finite float (madmom/audio/signal.py:395-438)."""
    if isinstance(signal, FramedSignal):
        return np.array([sound_pressure_level(frame) for frame in signal])
    rms = root_mean_square(signal)
    if p_ref is None:
        dtype = np.asarray(signal).dtype
        p_ref = float(np.iinfo(dtype).max) if np.issubdtype(dtype, np.integer) else 1.0
    with np.errstate(divide="ignore"):
        return float(np.nan_to_num(20.0 * np.log10(rms / p_ref)))


def write_wave_file(signal, filename, sample_rate=None):
    """Write the samples as a RIFF/WAVE file with scipy (madmom/io/audio.py:670-702)."""
    from scipy.io import wavfile
    if sample_rate is None:
        sample_rate = int(signal.sample_rate)
    wavfile.write(filename, rate=sample_rate, data=np.asarray(signal))
    return filename


def remix(signal, num_channels, channel=None):
    """Down-mix to mono (mean over channels cast back to the input dtype) or up-mix mono by repetition
    (madmom/audio/signal.py:170-223)."""
    if num_channels is None or num_channels == (1 if signal.ndim == 1 else signal.shape[1]):
        return signal
    if signal.ndim > 1 and num_channels == 1:
        if channel is None:
            return np.mean(signal, axis=-1).astype(signal.dtype)
        return signal[:, channel]
    if signal.ndim == 1:
        return np.tile(signal[:, np.newaxis], num_channels)
    raise NotImplementedError("Requested %d channels, but got %d channels and channel conversion is not "
                              "implemented." % (num_channels, signal.shape[1]))


class Signal(np.ndarray):
    """PCM samples with `sample_rate`, `start`, `stop` (madmom/audio/signal.py:600-732).  `data` is a numpy
    array or the name of a .wav file."""

    def __new__(cls, data, sample_rate=SAMPLE_RATE, num_channels=NUM_CHANNELS, channel=None, start=None,
                stop=None, norm=False, gain=0., dtype=None, **kwargs):
        if not isinstance(data, np.ndarray):
            data, sample_rate = load_wave_file(data, sample_rate=sample_rate, num_channels=num_channels,
                                               start=start, stop=stop, dtype=dtype)
        if not isinstance(data, Signal):
            data = np.asarray(data).view(cls)
            data.sample_rate = sample_rate
        if num_channels:
            data = remix(data, num_channels, channel).view(cls)
        if norm:
            data = normalize(data)
        if gain is not None and gain != 0:
            data = adjust_gain(data, gain)
        if sample_rate is not None and data.sample_rate is not None and sample_rate != data.sample_rate:
            raise NotImplementedError("re-sampling is outside the GPU front end")
        if data.sample_rate is None:
            data.sample_rate = sample_rate
        if start is not None:
            data.start = start
            data.stop = start + float(len(data)) / data.sample_rate
        return data

    def __array_finalize__(self, obj):
        if obj is None:
            return
        self.sample_rate = getattr(obj, "sample_rate", None)
        self.start = getattr(obj, "start", None)
        self.stop = getattr(obj, "stop", None)

    def __reduce__(self):
        state = np.ndarray.__reduce__(self)
        return state[0], state[1], state[2] + ((self.sample_rate, self.start, self.stop),)

    def __setstate__(self, state):
        np.ndarray.__setstate__(self, state[:-1])
        self.sample_rate, self.start, self.stop = state[-1]

    @property
    def num_samples(self):
        return len(self)

    @property
    def num_channels(self):
        return 1 if self.ndim == 1 else np.shape(self)[1]

    @property
    def length(self):
        if self.sample_rate is None:
            return None
        return float(self.num_samples) / self.sample_rate

    def write(self, filename):
        return write_wave_file(self, filename)

    def energy(self):
        return energy(self)

    def root_mean_square(self):
        return root_mean_square(self)

    rms = root_mean_square

    def sound_pressure_level(self, p_ref=None):
        return sound_pressure_level(self, p_ref)

    spl = sound_pressure_level


class SignalProcessor(Processor):
    """Load / wrap a signal (madmom/audio/signal.py:735-857)."""

    def __init__(self, sample_rate=SAMPLE_RATE, num_channels=NUM_CHANNELS, start=None, stop=None, norm=False,
                 gain=0., dtype=None, **kwargs):
        self.sample_rate = sample_rate
        self.num_channels = num_channels
        self.start = start
        self.stop = stop
        self.norm = norm
        self.gain = gain
        self.dtype = dtype

    def process(self, data, **kwargs):
        args = dict(sample_rate=self.sample_rate, num_channels=self.num_channels, start=self.start,
                    stop=self.stop, norm=self.norm, gain=self.gain, dtype=self.dtype)
        args.update(kwargs)
        return Signal(data, **args)


def signal_frame(signal, index, frame_size, hop_size, origin=0, pad=0):
    """Frame `index` of `signal` on the host (madmom/audio/signal.py:860-962, constant padding only)."""
    if pad != 0:
        raise NotImplementedError("only zero padding is supported")
    n = len(signal)
    start = frame_start(index, frame_size, hop_size, origin)
    stop = start + int(frame_size)
    if start >= 0 and stop <= n:
        return signal[start:stop]
    frame = np.zeros((int(frame_size),) + signal.shape[1:], dtype=signal.dtype).view(type(signal))
    if isinstance(signal, Signal):
        frame.sample_rate = signal.sample_rate
    lo, hi = max(start, 0), min(stop, n)
    if hi > lo:
        frame[lo - start:hi - start] = signal[lo:hi]
    return frame


class FramedSignal(object):
    """Lazy (num_frames, frame_size) view of a Signal (madmom/audio/signal.py:1103-1260)."""

    def __init__(self, signal, frame_size=FRAME_SIZE, hop_size=HOP_SIZE, fps=FPS, origin=ORIGIN,
                 end=END_OF_SIGNAL, num_frames=NUM_FRAMES, **kwargs):
        if not isinstance(signal, Signal):
            signal = Signal(signal, **kwargs)
        self.signal = signal
        if frame_size:
            self.frame_size = int(frame_size)
        if hop_size:
            self.hop_size = float(hop_size)
        if fps:
            self.hop_size = self.signal.sample_rate / float(fps)
        self.origin = resolve_origin(origin, frame_size)
        if num_frames is None:
            num_frames = _num_frames(len(self.signal), self.hop_size, end)
        self.num_frames = int(num_frames)
        self._device_pcm = None

    def __getstate__(self):
        state = dict(self.__dict__)
        state["_device_pcm"] = None
        return state

    def __getitem__(self, index):
        if isinstance(index, (int, np.integer)):
            if index < 0:
                index += self.num_frames
            if index < self.num_frames:
                return signal_frame(self.signal, index, self.frame_size, self.hop_size, self.origin)
            raise IndexError("end of signal reached")
        if isinstance(index, slice):
            start, stop, step = index.indices(self.num_frames)
            if step != 1:
                raise ValueError("only slices with a step size of 1 supported")
            # the slice is a new FramedSignal whose origin is shifted by `start` hops (truncated to int)
            return FramedSignal(self.signal, frame_size=self.frame_size, hop_size=self.hop_size,
                                origin=self.origin - self.hop_size * start, num_frames=stop - start)
        raise TypeError("frame indices must be slices or integers")

    def __len__(self):
        return self.num_frames

    @property
    def frame_rate(self):
        if self.signal.sample_rate is None:
            return None
        return float(self.signal.sample_rate) / self.hop_size

    fps = frame_rate

    @property
    def overlap_factor(self):
        return 1.0 - self.hop_size / self.frame_size

    @property
    def shape(self):
        shape = self.num_frames, self.frame_size
        if self.signal.num_channels != 1:
            shape += (self.signal.num_channels,)
        return shape

    @property
    def ndim(self):
        return len(self.shape)

    def device_pcm(self, device):
        """The signal as a contiguous CUDA tensor (uploaded once, shared by every stage / frame size)."""
        import torch
        if self._device_pcm is None or self._device_pcm.device.index != device:
            host = np.ascontiguousarray(np.asarray(self.signal))
            self._device_pcm = torch.from_numpy(host).to("cuda:%d" % device, non_blocking=False)
        return self._device_pcm


class FramedSignalProcessor(Processor):
    """Slice a Signal into frames (madmom/audio/signal.py:1263-1358)."""

    def __init__(self, frame_size=FRAME_SIZE, hop_size=HOP_SIZE, fps=FPS, origin=ORIGIN, end=END_OF_SIGNAL,
                 num_frames=NUM_FRAMES, **kwargs):
        self.frame_size = frame_size
        self.hop_size = hop_size
        self.fps = fps
        self.origin = origin
        self.end = end
        self.num_frames = num_frames

    def process(self, data, **kwargs):
        args = dict(frame_size=self.frame_size, hop_size=self.hop_size, fps=self.fps, origin=self.origin,
                    end=self.end, num_frames=self.num_frames)
        args.update(kwargs)
        return FramedSignal(data, **args)
