"""Filterbanks of the spectral front end (host side, plan constants).

Mirror of the part of madmom/audio/filters.py the hot path uses: `log_frequencies`, `mel_frequencies`,
`frequencies2bins`, `bins2frequencies`, `Filterbank`, `LogarithmicFilterbank`, `MelFilterbank`
(madmom/audio/filters.py:21-78, 186-223, 348-410, 692-873, 1035-1092, 1153-1237).

A filterbank is built once per plan on the host (exactly as the reference does, float64 geometry ->
float32 weights) and handed to the CUDA library as a *band table*: one contiguous run of weights per
band (`Filterbank.band_table()`), which is what the fused kernel contracts against the magnitudes in
shared memory.  The dense (num_bins, num_bands) float32 ndarray view is kept because callers of the
reference index and multiply it directly (e.g. complex_flux reads the non-zero bins of every band).
"""
import numpy as np

FILTER_DTYPE = np.float32
A4 = 440.


# ---------------------------------------------------------------------------------------------- scales
def hz2mel(f):
    """madmom/audio/filters.py:21-36"""
    return 1127.01048 * np.log(np.asarray(f) / 700. + 1.)


def mel2hz(m):
    """madmom/audio/filters.py:39-54"""
    return 700. * (np.exp(np.asarray(m) / 1127.01048) - 1.)


def mel_frequencies(num_bands, fmin, fmax):
    """`num_bands` frequencies equally spaced on the Mel scale (madmom/audio/filters.py:57-78)."""
    return mel2hz(np.linspace(hz2mel(fmin), hz2mel(fmax), num_bands))


def log_frequencies(bands_per_octave, fmin, fmax, fref=A4):
    """Frequencies fref * 2^(k / bands_per_octave) inside [fmin, fmax] (madmom/audio/filters.py:186-223)."""
    k_lo = np.floor(np.log2(float(fmin) / fref) * bands_per_octave)
    k_hi = np.ceil(np.log2(float(fmax) / fref) * bands_per_octave)
    freqs = fref * 2. ** (np.arange(k_lo, k_hi) / float(bands_per_octave))
    lo = np.searchsorted(freqs, fmin)
    hi = np.searchsorted(freqs, fmax, "right")
    return freqs[lo:hi]


def frequencies2bins(frequencies, bin_frequencies, unique_bins=False):
    """Index of the closest bin for every frequency; a tie goes to the lower bin
    (madmom/audio/filters.py:348-388)."""
    frequencies = np.asarray(frequencies)
    bin_frequencies = np.asarray(bin_frequencies)
    right = np.clip(bin_frequencies.searchsorted(frequencies), 1, len(bin_frequencies) - 1)
    d_left = frequencies - bin_frequencies[right - 1]
    d_right = bin_frequencies[right] - frequencies
    bins = right - (d_left < d_right)
    return np.unique(bins) if unique_bins else bins


def bins2frequencies(bins, bin_frequencies):
    """madmom/audio/filters.py:391-410"""
    return np.asarray(bin_frequencies, dtype=float)[np.asarray(bins)]


# ---------------------------------------------------------------------------------------------- bands
def _triangle(start, center, stop, norm):
    """float32 weights of one triangle on bins [start, stop) peaking at `center`
    (TriangularFilter.__new__ + Filter.__new__, madmom/audio/filters.py:440-458, 532-555): float64
    linspace ramps, cast to float32, then normalised by the float32 sum."""
    if not start <= center < stop:
        raise ValueError("`center` must be between `start` and `stop`")
    up, down = center - start, stop - center
    w = np.empty(up + down)
    w[:up] = np.linspace(0, 1, up, endpoint=False)
    w[up:] = np.linspace(1, 0, down, endpoint=False)
    w = w.astype(FILTER_DTYPE)
    if norm:
        w /= np.sum(w)
    return w


def overlapping_triangles(bins, norm):
    """[(start, weights)] of overlapping triangles over consecutive bin triples; a triple spanning less
    than two bins degenerates to the single bin `start` with weight 1
    (TriangularFilter.band_bins / Filter.filters, madmom/audio/filters.py:476-502, 562-605)."""
    bins = [int(b) for b in bins]
    if len(bins) < 3:
        raise ValueError("not enough bins to create a TriangularFilter")
    bands = []
    for start, center, stop in zip(bins[:-2], bins[1:-1], bins[2:]):
        if stop - start < 2:
            center, stop = start, start + 1
        bands.append((start, _triangle(start, center, stop, norm)))
    return bands


class Filterbank(np.ndarray):
    """Dense (num_bins, num_bands) float32 filterbank with `bin_frequencies`
    (madmom/audio/filters.py:692-873)."""

    def __new__(cls, data, bin_frequencies):
        if not (isinstance(data, np.ndarray) and data.ndim == 2):
            raise TypeError("wrong input data for Filterbank, must be a 2D np.ndarray")
        obj = np.asarray(data, dtype=FILTER_DTYPE).view(cls)
        if len(bin_frequencies) != obj.shape[0]:
            raise ValueError("`bin_frequencies` must have the same length as the first dimension of `data`.")
        obj.bin_frequencies = np.asarray(bin_frequencies, dtype=float)
        return obj

    def __array_finalize__(self, obj):
        if obj is None:
            return
        self.bin_frequencies = getattr(obj, "bin_frequencies", None)

    def __reduce__(self):
        state = np.ndarray.__reduce__(self)
        return state[0], state[1], state[2] + (self.__dict__,)

    def __setstate__(self, state):
        np.ndarray.__setstate__(self, state[:-1])
        self.__dict__.update(state[-1])

    @classmethod
    def from_filters(cls, filters, bin_frequencies):
        """Place [(start, weights)] bands into a zero matrix, keeping the element-wise maximum where a
        band is placed twice, truncating at the edges (madmom/audio/filters.py:746-821)."""
        num_bins = len(bin_frequencies)
        dense = np.zeros((num_bins, len(filters)))
        for b, (start, w) in enumerate(filters):
            stop = start + len(w)
            lo, hi = max(start, 0), min(stop, num_bins)
            if hi > lo:
                dense[lo:hi, b] = np.maximum(dense[lo:hi, b], w[lo - start:hi - start])
        return Filterbank.__new__(cls, dense, bin_frequencies)

    # -------------------------------------------------------------------------------- band table
    def band_table(self):
        """(start[B], length[B], weights[nnz]) int32/int32/float32: per band the run first..last non-zero
        bin (interior zeros kept).  This is the filterbank layout of include/madmom_b200.h."""
        dense = np.asarray(self)
        starts = np.zeros(dense.shape[1], np.int32)
        lens = np.zeros(dense.shape[1], np.int32)
        runs = []
        for b in range(dense.shape[1]):
            nz = np.flatnonzero(dense[:, b])
            if nz.size:
                starts[b], lens[b] = nz[0], nz[-1] + 1 - nz[0]
                runs.append(dense[nz[0]:nz[-1] + 1, b])
        w = np.concatenate(runs).astype(FILTER_DTYPE) if runs else np.zeros(0, FILTER_DTYPE)
        return starts, lens, np.ascontiguousarray(w)

    @property
    def num_bins(self):
        return self.shape[0]

    @property
    def num_bands(self):
        return self.shape[1]

    def _band_edges(self):
        s, l, _ = self.band_table()
        return s.astype(int), (s + l - 1).astype(int)

    @property
    def corner_frequencies(self):
        lo, hi = self._band_edges()
        return bins2frequencies(np.stack((lo, hi), axis=1), self.bin_frequencies)

    @property
    def center_frequencies(self):
        """Peak bin of every band, or the middle bin of a flat band (madmom/audio/filters.py:845-862)."""
        lo, hi = self._band_edges()
        dense = np.asarray(self)
        centers = []
        for b, (a, z) in enumerate(zip(lo, hi)):
            if dense[a, b] == dense[z, b]:
                centers.append(int(a + (z - a) / 2.))
            else:
                centers.append(a + int(np.argmax(dense[a:z, b])))
        return bins2frequencies(centers, self.bin_frequencies)

    @property
    def fmin(self):
        return self.bin_frequencies[np.nonzero(np.asarray(self))[0][0]]

    @property
    def fmax(self):
        return self.bin_frequencies[np.nonzero(np.asarray(self))[0][-1]]


class LogarithmicFilterbank(Filterbank):
    """Triangular filters on a logarithmic frequency axis (madmom/audio/filters.py:1153-1237)."""
    NUM_BANDS_PER_OCTAVE = 12
    FMIN = 30.
    FMAX = 17000.
    NORM_FILTERS = True
    UNIQUE_FILTERS = True

    def __new__(cls, bin_frequencies, num_bands=NUM_BANDS_PER_OCTAVE, fmin=FMIN, fmax=FMAX, fref=A4,
                norm_filters=NORM_FILTERS, unique_filters=UNIQUE_FILTERS, bands_per_octave=True):
        if not bands_per_octave:
            raise NotImplementedError("please implement `num_bands` with `bands_per_octave` set to 'False' "
                                      "for LogarithmicFilterbank")
        freqs = log_frequencies(num_bands, fmin, fmax, fref)
        bins = frequencies2bins(freqs, bin_frequencies, unique_bins=unique_filters)
        obj = cls.from_filters(overlapping_triangles(bins, norm_filters), bin_frequencies)
        obj.fref = fref
        obj.num_bands_per_octave = num_bands
        return obj

    def __array_finalize__(self, obj):
        if obj is None:
            return
        self.bin_frequencies = getattr(obj, "bin_frequencies", None)
        self.num_bands_per_octave = getattr(obj, "num_bands_per_octave", self.NUM_BANDS_PER_OCTAVE)
        self.fref = getattr(obj, "fref", A4)


LogFilterbank = LogarithmicFilterbank


class MelFilterbank(Filterbank):
    """Triangular filters on the Mel scale (madmom/audio/filters.py:1035-1092)."""
    NUM_BANDS = 40
    FMIN = 20.
    FMAX = 17000.
    NORM_FILTERS = True
    UNIQUE_FILTERS = True

    def __new__(cls, bin_frequencies, num_bands=NUM_BANDS, fmin=FMIN, fmax=FMAX, norm_filters=NORM_FILTERS,
                unique_filters=UNIQUE_FILTERS, **kwargs):
        freqs = mel_frequencies(num_bands + 2, fmin, fmax)      # two more: the outer edges
        bins = frequencies2bins(freqs, bin_frequencies, unique_bins=unique_filters)
        return cls.from_filters(overlapping_triangles(bins, norm_filters), bin_frequencies)
