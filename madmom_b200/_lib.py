"""ctypes binding of libmadmom_b200.so (include/madmom_b200.h) and the thin `FrontEndPlan` wrapper used by
the processor classes.  There is no fallback: if the library cannot be loaded every compute call raises."""
import ctypes
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIBPATH = os.path.join(_HERE, "lib", "libmadmom_b200.so")

MMB_OK, MMB_E_INVALID, MMB_E_UNSUPPORTED, MMB_E_CUDA, MMB_E_NOMEM = 0, -1, -2, -3, -4
PCM_I16, PCM_F32 = 0, 1
LOG_NONE, LOG_LOG10, LOG_MULADD, LOG_LN, LOG_LOG2 = 0, 1, 2, 3, 4

c_i32, c_i64, c_f32, c_f64, c_vp = ctypes.c_int32, ctypes.c_int64, ctypes.c_float, ctypes.c_double, ctypes.c_void_p


class Config(ctypes.Structure):
    _fields_ = [("frame_size", c_i32), ("pcm_dtype", c_i32), ("hop_size", c_f64), ("origin", c_i32),
                ("circular_shift", c_i32), ("fft_window", ctypes.POINTER(c_f64)),
                ("num_bands", c_i32), ("band_start", ctypes.POINTER(c_i32)), ("band_len", ctypes.POINTER(c_i32)),
                ("band_weights", ctypes.POINTER(c_f32)), ("log_mode", c_i32), ("mul", c_f32), ("add", c_f32),
                ("diff_frames", c_i32), ("diff_max_bins", c_i32), ("positive_diffs", c_i32),
                ("chunk_frames", c_i32), ("device", c_i32), ("include_nyquist", c_i32)]


class Track(ctypes.Structure):
    _fields_ = [("pcm_offset", c_i64), ("pcm_len", c_i64), ("sample_base", c_i64), ("frame_first", c_i64),
                ("frame_count", c_i64), ("out_row", c_i64)]


TRACK_DTYPE = np.dtype([("pcm_offset", "<i8"), ("pcm_len", "<i8"), ("sample_base", "<i8"), ("frame_first", "<i8"),
                        ("frame_count", "<i8"), ("out_row", "<i8")])
_UNIT_FIELD = {name: i for i, name in enumerate(TRACK_DTYPE.names)}


class Outputs(ctypes.Structure):
    _fields_ = [("stft", c_vp), ("ld_stft", c_i64), ("spec", c_vp), ("ld_spec", c_i64), ("filt", c_vp),
                ("ld_filt", c_i64), ("log", c_vp), ("ld_log", c_i64), ("diff", c_vp), ("ld_diff", c_i64),
                ("odf", c_vp), ("lgd", c_vp), ("ld_lgd", c_i64)]


class PlanInfo(ctypes.Structure):
    _fields_ = [(n, c_i32) for n in ("frame_size", "num_bins", "num_bands", "chunk_frames", "smem_bytes",
                                     "ctas_per_sm", "num_sms", "frames_per_round", "warps")]


EXPORTS = {
    "mmb_version": (c_i32, []),
    "mmb_last_error": (ctypes.c_char_p, []),
    "mmb_num_frames": (c_i64, [c_i64, c_f64, c_i32]),
    "mmb_frame_start": (c_i64, [c_i64, c_i32, c_f64, c_i32]),
    "mmb_plan_create": (c_i32, [ctypes.POINTER(Config), ctypes.POINTER(c_vp)]),
    "mmb_plan_destroy": (c_i32, [c_vp]),
    "mmb_plan_get_info": (c_i32, [c_vp, ctypes.POINTER(PlanInfo)]),
    "mmb_layout_create": (c_i32, [c_vp, c_vp, c_i32, ctypes.POINTER(c_vp)]),
    "mmb_layout_destroy": (c_i32, [c_vp]),
    "mmb_run": (c_i32, [c_vp, c_vp, c_vp, ctypes.POINTER(Outputs), c_vp, ctypes.POINTER(c_i32)]),
    "mmb_run_gather": (c_i32, [c_vp, c_vp, c_vp, ctypes.POINTER(Outputs), c_i32, c_vp, ctypes.POINTER(c_i32)]),
    "mmb_run_multicast": (c_i32, [c_vp, c_vp, c_vp, ctypes.POINTER(Outputs), c_vp, ctypes.POINTER(c_i32)]),
    "mmb_run_host": (c_i32, [c_vp, c_vp, c_i32, c_vp, ctypes.POINTER(Outputs), ctypes.POINTER(c_i32)]),
    "mmb_run_host_interleaved": (c_i32, [c_vp, c_vp, c_i32, c_vp, c_i32, c_i32, ctypes.POINTER(Outputs),
                                         ctypes.POINTER(c_i32)]),
    "mmb_remix_mono": (c_i32, [c_vp, c_i32, c_i64, c_i32, c_i32, c_vp, c_vp]),
    "mmb_host_alloc": (c_i32, [ctypes.POINTER(c_vp), c_i64]),
    "mmb_host_free": (c_i32, [c_vp]),
    "mmb_measure_fp32_peak": (c_i32, [c_i32, ctypes.POINTER(c_f64)]),
    "mmb_rate_probe": (c_i32, [c_i32, c_i32, ctypes.POINTER(c_f64)]),
    "mmb_fft_probe": (c_i32, [c_i32, c_i32, ctypes.POINTER(c_f64)]),
    "mmb_pipeline_probe": (c_i32, [c_i32, c_i32, ctypes.POINTER(c_f64)]),
    "mmb_matrix_diff": (c_i32, [c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_i32, c_i64, c_i32, c_i32, c_i32, c_vp]),
    "mmb_complex_flux": (c_i32, [c_vp, c_i64, c_vp, c_i64, c_vp, c_vp, c_i32, c_i32, c_i64, c_i64, c_vp, c_i32, c_i32,
                                 c_vp, c_vp]),
    "mmb_spectral_odf": (c_i32, [c_i32, c_vp, c_i64, c_vp, c_i64, c_i64, c_i32, c_i64, c_vp, c_i32, c_i32, c_f64,
                                 c_vp, c_vp]),
    "mmb_peak_picking": (c_i32, [c_vp, c_vp, c_i32, c_i64, c_f64, c_i32, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp,
                                 c_vp, c_vp]),
    "mmb_phase_lgd": (c_i32, [c_vp, c_i64, c_vp, c_i64, c_vp, c_i64, c_i64, c_i32, c_vp]),
    "mmb_peak_picking_static": (c_i32, [c_vp, c_vp, c_i32, c_i64, c_f64, c_i32, c_vp, c_i32, c_i32, c_i32, c_i32,
                                        c_vp, c_i32, c_vp, c_vp]),
}

_lib = None
_lock = threading.Lock()
launch_count = 0          # kernels launched through this binding (bench.py reports it as gpu_launches)


class LibraryError(RuntimeError):
    pass


def load():
    """Load the shared library (once).  Raises LibraryError if it is missing -- there is no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIBPATH):
            raise LibraryError("%s not found: build it with `python -m madmom_b200.build` "
                               "(madmom_b200 has no CPU fallback)" % LIBPATH)
        lib = ctypes.CDLL(LIBPATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def _check(rc):
    if rc == MMB_OK:
        return
    msg = load().mmb_last_error().decode("utf-8", "replace")
    if rc == MMB_E_UNSUPPORTED:
        raise NotImplementedError(msg)
    if rc == MMB_E_INVALID:
        raise ValueError(msg)
    if rc == MMB_E_NOMEM:
        raise MemoryError(msg)
    raise LibraryError("libmadmom_b200: %s (code %d)" % (msg, rc))


def num_frames(num_samples, hop_size, end="normal"):
    if end not in ("normal", "extend"):
        raise ValueError("end of signal handling '%s' unknown" % end)
    return int(load().mmb_num_frames(int(num_samples), float(hop_size), 1 if end == "extend" else 0))


def frame_start(index, frame_size, hop_size, origin=0):
    return int(load().mmb_frame_start(int(index), int(frame_size), float(hop_size), int(origin)))


def measure_fp32_peak(device=0):
    """Measured FP32 FMA peak of the device in TFLOP/s (mmb_measure_fp32_peak)."""
    v = c_f64(0.0)
    _check(load().mmb_measure_fp32_peak(int(device), ctypes.byref(v)))
    return v.value


def fft_only_frames_per_s(device=0, frame_size=2048):
    """Frames per second of the fused kernel's warp FFT alone (mmb_fft_probe): the ceiling it is judged against."""
    v = c_f64(0.0)
    _check(load().mmb_fft_probe(int(device), int(frame_size), ctypes.byref(v)))
    return v.value


def compact_filterbank(fb):
    """(K, B) dense float32 filterbank -> (start[B], len[B], weights[nnz]): one run min..max nonzero per band."""
    fb = np.asarray(fb, dtype=np.float32)
    starts, lens, ws = [], [], []
    for b in range(fb.shape[1]):
        nz = np.nonzero(fb[:, b])[0]
        if len(nz) == 0:
            starts.append(0), lens.append(0)
            continue
        s, e = int(nz[0]), int(nz[-1]) + 1
        starts.append(s), lens.append(e - s), ws.append(fb[s:e, b])
    w = np.concatenate(ws).astype(np.float32) if ws else np.zeros(0, np.float32)
    return np.asarray(starts, np.int32), np.asarray(lens, np.int32), np.ascontiguousarray(w)


def make_units(num_samples, frame_counts, pcm_offsets=None, sample_base=None, frame_first=None):
    """Unit table for whole tracks laid back to back (or explicit offsets)."""
    num_samples = np.asarray(num_samples, dtype=np.int64)
    frame_counts = np.asarray(frame_counts, dtype=np.int64)
    n = len(num_samples)
    u = np.zeros(n, dtype=TRACK_DTYPE)
    u["pcm_len"] = num_samples
    u["pcm_offset"] = np.concatenate(([0], np.cumsum(num_samples)[:-1])) if pcm_offsets is None else pcm_offsets
    u["sample_base"] = 0 if sample_base is None else sample_base
    u["frame_first"] = 0 if frame_first is None else frame_first
    u["frame_count"] = frame_counts
    u["out_row"] = np.concatenate(([0], np.cumsum(frame_counts)[:-1]))
    return u


OUTPUT_NAMES = ("stft", "spec", "filt", "log", "diff", "odf", "lgd")


class FrontEndPlan(object):
    """One fused-kernel recipe (mmb_plan).  `filterbank` is a dense (K, B) array or None."""

    def __init__(self, frame_size, hop_size, fft_window, pcm_dtype=np.int16, origin=0, circular_shift=False,
                 filterbank=None, log_mode=LOG_NONE, mul=1.0, add=0.0, diff_frames=0, diff_max_bins=0,
                 positive_diffs=False, chunk_frames=0, device=0, include_nyquist=False):
        lib = load()
        self._lib = lib
        self._handle = c_vp()
        self._layouts = {}
        self._layout_lock = threading.Lock()
        self.frame_size, self.hop_size, self.origin = int(frame_size), float(hop_size), int(origin)
        self.num_bins = self.frame_size // 2 + (1 if include_nyquist else 0)
        self.device = int(device)
        pcm_dtype = np.dtype(pcm_dtype)
        if pcm_dtype == np.int16:
            code = PCM_I16
        elif pcm_dtype == np.float32:
            code = PCM_F32
        else:
            raise NotImplementedError("PCM dtype %s not supported on the GPU path (int16 or float32)" % pcm_dtype)
        self.pcm_dtype = pcm_dtype
        win = np.ascontiguousarray(np.asarray(fft_window, dtype=np.float64))
        if win.shape != (self.frame_size,):
            raise ValueError("fft_window must have frame_size elements")
        cfg = Config()
        cfg.frame_size, cfg.pcm_dtype, cfg.hop_size, cfg.origin = self.frame_size, code, self.hop_size, self.origin
        cfg.circular_shift = int(bool(circular_shift))
        cfg.include_nyquist = int(bool(include_nyquist))
        cfg.fft_window = win.ctypes.data_as(ctypes.POINTER(c_f64))
        keep = [win]
        if filterbank is not None:
            s, l, w = compact_filterbank(filterbank)
            keep += [s, l, w]
            cfg.num_bands = len(s)
            cfg.band_start = s.ctypes.data_as(ctypes.POINTER(c_i32))
            cfg.band_len = l.ctypes.data_as(ctypes.POINTER(c_i32))
            cfg.band_weights = w.ctypes.data_as(ctypes.POINTER(c_f32))
        self.num_bands = int(cfg.num_bands)
        cfg.log_mode, cfg.mul, cfg.add = int(log_mode), float(mul), float(add)
        cfg.diff_frames, cfg.diff_max_bins = int(diff_frames), int(diff_max_bins or 0)
        cfg.positive_diffs, cfg.chunk_frames, cfg.device = int(bool(positive_diffs)), int(chunk_frames), self.device
        self.diff_frames = int(diff_frames)
        _check(lib.mmb_plan_create(ctypes.byref(cfg), ctypes.byref(self._handle)))
        info = PlanInfo()
        _check(lib.mmb_plan_get_info(self._handle, ctypes.byref(info)))
        self.info = {n: getattr(info, n) for n, _ in PlanInfo._fields_}

    def __del__(self):
        try:
            for h in self._layouts.values():
                self._lib.mmb_layout_destroy(h)
            if self._handle:
                self._lib.mmb_plan_destroy(self._handle)
        except Exception:
            pass

    def width(self, name):
        return {"stft": self.num_bins, "spec": self.num_bins, "lgd": self.num_bins, "odf": 1}.get(name, self.num_bands)

    def _layout(self, units):
        """Device copy of a unit table, cached by content (least recently used of 64 kept).  Destroying a layout
        frees device memory, which waits for the kernels that read it, so an evicted layout is never pulled from
        under a launch already enqueued; the lock keeps two host threads from evicting under each other."""
        key = units.tobytes()
        with self._layout_lock:
            h = self._layouts.pop(key, None)
            if h is None:
                while len(self._layouts) >= 64:
                    old = self._layouts.pop(next(iter(self._layouts)))
                    self._lib.mmb_layout_destroy(old)
                h = c_vp()
                _check(self._lib.mmb_layout_create(self._handle, units.ctypes.data_as(c_vp), len(units),
                                                   ctypes.byref(h)))
            self._layouts[key] = h                       # (re-)inserted last: most recently used
        return h

    @staticmethod
    def _check_extents(units, pcm_elems, out_rows):
        """The unit table must stay inside the tensors handed in: the kernel trusts it."""
        if len(units) == 0:
            return
        if len(units) == 1:                                  # one file at a time: no numpy temporaries (latency path)
            u = units[0].item()                              # one tuple of Python ints
            need_pcm = u[_UNIT_FIELD["pcm_offset"]] + u[_UNIT_FIELD["pcm_len"]]
            need_rows = u[_UNIT_FIELD["out_row"]] + u[_UNIT_FIELD["frame_count"]]
            if need_pcm > pcm_elems:
                raise ValueError("unit table reads %d PCM samples, the pcm tensor has %d" % (need_pcm, pcm_elems))
            for name, rows in out_rows.items():
                if rows < need_rows:
                    raise ValueError("output %r has %d rows, the unit table writes %d" % (name, rows, need_rows))
            return
        need_pcm = int((units["pcm_offset"] + units["pcm_len"]).max())
        if need_pcm > pcm_elems:
            raise ValueError("unit table reads %d PCM samples, the pcm tensor has %d" % (need_pcm, pcm_elems))
        need_rows = int((units["out_row"] + units["frame_count"]).max())
        for name, rows in out_rows.items():
            if rows < need_rows:
                raise ValueError("output %r has %d rows, the unit table writes %d" % (name, rows, need_rows))

    @staticmethod
    def _fill(outs, ptrs):
        o = Outputs()
        for name, (ptr, ld) in ptrs.items():
            setattr(o, name, ptr)
            if name != "odf":
                setattr(o, "ld_" + name, ld)
        return o

    def run(self, pcm, units, outputs, stream=None):
        """Device path.  pcm: CUDA tensor; outputs: {name: CUDA tensor (rows, >= width) or column view}.
        Asynchronous on the current torch stream."""
        global launch_count
        import torch
        units = np.ascontiguousarray(units, dtype=TRACK_DTYPE)
        if not pcm.is_cuda or pcm.device.index != self.device:
            raise ValueError("pcm must live on cuda:%d" % self.device)
        want = torch.int16 if self.pcm_dtype == np.int16 else torch.float32
        if pcm.dtype != want or not pcm.is_contiguous():
            raise ValueError("pcm must be a contiguous %s tensor" % want)
        ptrs = {}
        for name, t in outputs.items():
            if name not in OUTPUT_NAMES:
                raise ValueError("unknown output %r" % name)
            if self.num_bands == 0 and name in ("filt", "log", "diff", "odf"):
                raise NotImplementedError("output %r needs a filterbank in the plan" % name)
            if t.dtype != (torch.complex64 if name == "stft" else torch.float32) or not t.is_cuda:
                raise ValueError("output %s has the wrong dtype/device" % name)
            if name == "odf":
                if not t.is_contiguous():
                    raise ValueError("odf output must be contiguous")
                ptrs[name] = (t.data_ptr(), 1)
            else:
                if t.dim() != 2 or t.stride(1) != 1 or t.shape[1] != self.width(name):
                    raise ValueError("output %s must be (rows, %d) with unit column stride" % (name, self.width(name)))
                if t.stride(0) < self.width(name) and t.shape[0] > 1:
                    raise ValueError("output %s: row stride smaller than the row width" % name)
                ptrs[name] = (t.data_ptr(), t.stride(0))
        self._check_extents(units, pcm.numel(), {name: t.shape[0] for name, t in outputs.items()})
        o = self._fill(outputs, ptrs)
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        n = c_i32(0)
        _check(self._lib.mmb_run(self._handle, self._layout(units), pcm.data_ptr(), ctypes.byref(o), stream,
                                 ctypes.byref(n)))
        launch_count += n.value
        return n.value

    def run_gather(self, pcm, units, destinations, stream=None, multicast=False):
        """Device path fused with an all-gather (mmb_run_gather): `destinations` is a list of {name: (pointer,
        leading dimension)} with name in log / diff / odf; entry 0 is this GPU's memory, the others are the peers'
        buffers mapped into this process (torch symmetric memory), every pointer already offset to this rank's
        rows.  Each row is stored to all destinations by the kernel.  multicast=True: `destinations` is ONE entry
        of NVLS multicast addresses (mmb_run_multicast: one multimem.st per value, the NVSwitch replicates it).
        Asynchronous on the current torch stream."""
        global launch_count
        import torch
        units = np.ascontiguousarray(units, dtype=TRACK_DTYPE)
        if not pcm.is_cuda or pcm.device.index != self.device or not pcm.is_contiguous():
            raise ValueError("pcm must be a contiguous tensor on cuda:%d" % self.device)
        arr = (Outputs * len(destinations))()
        for i, dst in enumerate(destinations):
            for name, (ptr, ld) in dst.items():
                if name not in ("log", "diff", "odf"):
                    raise ValueError("run_gather stores the outputs log, diff and odf only, not %r" % name)
                setattr(arr[i], name, ptr)
                if name != "odf":
                    setattr(arr[i], "ld_" + name, ld)
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        n = c_i32(0)
        if multicast:
            if len(destinations) != 1:
                raise ValueError("a multicast launch has one destination: the multicast addresses")
            _check(self._lib.mmb_run_multicast(self._handle, self._layout(units), pcm.data_ptr(), arr, stream,
                                               ctypes.byref(n)))
        else:
            _check(self._lib.mmb_run_gather(self._handle, self._layout(units), pcm.data_ptr(), arr,
                                            len(destinations), stream, ctypes.byref(n)))
        launch_count += n.value
        return n.value

    def run_host(self, pcm, units, outputs, num_channels=1, channel=None):
        """Host path (mmb_run_host).  pcm: contiguous numpy array or CPU tensor; outputs: {name: numpy array
        or CPU tensor}; returns when the outputs are complete.  num_channels > 1: `pcm` holds interleaved sample
        frames (unit offsets / lengths count frames) and is mixed down on the device (`channel` None: mean over the
        channels as madmom.audio.signal.remix does, else that channel)."""
        global launch_count
        units = np.ascontiguousarray(units, dtype=TRACK_DTYPE)

        def addr(a):
            return a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data

        def strides(a):
            if hasattr(a, "data_ptr"):
                return a.stride(0) if a.dim() == 2 else 1
            return a.strides[0] // a.itemsize if a.ndim == 2 else 1
        ptrs = {name: (addr(a), strides(a)) for name, a in outputs.items()}
        o = self._fill(outputs, ptrs)
        n = c_i32(0)
        if num_channels == 1:
            _check(self._lib.mmb_run_host(self._handle, units.ctypes.data_as(c_vp), len(units), addr(pcm),
                                          ctypes.byref(o), ctypes.byref(n)))
        else:
            _check(self._lib.mmb_run_host_interleaved(self._handle, units.ctypes.data_as(c_vp), len(units),
                                                      addr(pcm), int(num_channels),
                                                      -1 if channel is None else int(channel), ctypes.byref(o),
                                                      ctypes.byref(n)))
        launch_count += n.value
        return n.value


def _stream(device):
    import torch
    return torch.cuda.current_stream(device).cuda_stream


def matrix_diff(matrix, diff_frames, diff_max_bins=0, positive_diffs=False, rows_per_signal=0, want_diff=True,
                want_odf=False):
    """mmb_matrix_diff on a (rows, cols) float32 CUDA tensor.  Returns (diff or None, odf or None)."""
    global launch_count
    import torch
    if not matrix.is_cuda or matrix.dtype != torch.float32 or matrix.dim() != 2 or matrix.stride(1) != 1:
        raise ValueError("matrix must be a 2-D float32 CUDA tensor with unit column stride")
    rows, cols = matrix.shape
    diff = torch.empty((rows, cols), dtype=torch.float32, device=matrix.device) if want_diff else None
    odf = torch.empty((rows,), dtype=torch.float32, device=matrix.device) if want_odf else None
    rc = load().mmb_matrix_diff(matrix.data_ptr(), matrix.stride(0), diff.data_ptr() if want_diff else None,
                                cols, odf.data_ptr() if want_odf else None, rows, cols, int(rows_per_signal),
                                int(diff_frames), int(diff_max_bins or 0), int(bool(positive_diffs)),
                                _stream(matrix.device.index))
    if rc != MMB_OK:
        raise ValueError("mmb_matrix_diff: invalid argument") if rc == MMB_E_INVALID else LibraryError("mmb_matrix_diff failed")
    launch_count += 1 if rows else 0
    return diff, odf


def peak_picking(act, threshold, smooth=0, pre_avg=0, post_avg=0, pre_max=1, post_max=1, seg_offsets=None):
    """mmb_peak_picking.  act: 1-D float32 CUDA tensor; seg_offsets: None or int64 CUDA tensor [n_seg + 1].
    Returns (peaks int32 tensor [n], counts int32 tensor [n_seg]) on the device."""
    global launch_count
    import torch
    if not act.is_cuda or act.dtype != torch.float32 or act.dim() != 1 or not act.is_contiguous():
        raise ValueError("activations must be a contiguous 1-D float32 CUDA tensor")
    n = act.shape[0]
    n_seg = 1 if seg_offsets is None else int(seg_offsets.shape[0]) - 1
    peaks = torch.empty((max(n, 1),), dtype=torch.int32, device=act.device)
    counts = torch.zeros((n_seg,), dtype=torch.int32, device=act.device)
    kern = None
    if smooth and smooth > 1:
        kern = torch.from_numpy(np.hamming(int(smooth))).to(act.device)
    rc = load().mmb_peak_picking(act.data_ptr(), None if seg_offsets is None else seg_offsets.data_ptr(), n_seg, n,
                                 float(threshold), int(smooth or 0), None if kern is None else kern.data_ptr(),
                                 int(pre_avg), int(post_avg), int(pre_max), int(post_max), peaks.data_ptr(),
                                 counts.data_ptr(), _stream(act.device.index))
    if rc != MMB_OK:
        raise ValueError("mmb_peak_picking: invalid argument") if rc == MMB_E_INVALID else LibraryError("mmb_peak_picking failed")
    launch_count += 2 if n else 1
    return peaks, counts


STATIC_PEAK_MAX_FRAMES = 1 << 22


def peak_picking_static(act, threshold, peaks, counts, smooth=0, smooth_kernel=None, pre_avg=0, post_avg=0,
                        pre_max=1, post_max=1, seg_offsets=None):
    """mmb_peak_picking_static: peak picking that allocates nothing.  Every buffer is the caller's: `act` float32
    CUDA vector, `peaks` int32 CUDA tensor (n_segments, peaks_per_segment), `counts` int32 CUDA tensor (n_segments,),
    `smooth_kernel` float64 CUDA vector of length `smooth` when smooth > 1, `seg_offsets` int64 CUDA tensor or None.
    The flags live in a static device array of the library (at most STATIC_PEAK_MAX_FRAMES frames per call; calls
    that overlap on the GPU race).  Asynchronous on the current stream."""
    global launch_count
    import torch
    if not act.is_cuda or act.dtype != torch.float32 or act.dim() != 1 or not act.is_contiguous():
        raise ValueError("activations must be a contiguous 1-D float32 CUDA tensor")
    n_seg = 1 if seg_offsets is None else int(seg_offsets.shape[0]) - 1
    if (peaks.dtype != torch.int32 or peaks.dim() != 2 or peaks.shape[0] != n_seg or not peaks.is_contiguous() or
            counts.dtype != torch.int32 or counts.shape != (n_seg,) or not counts.is_contiguous()):
        raise ValueError("peaks must be a contiguous (n_segments, capacity) and counts a (n_segments,) int32 tensor")
    _check(load().mmb_peak_picking_static(act.data_ptr(), None if seg_offsets is None else seg_offsets.data_ptr(),
                                          n_seg, act.shape[0], float(threshold), int(smooth or 0),
                                          None if smooth_kernel is None else smooth_kernel.data_ptr(), int(pre_avg),
                                          int(post_avg), int(pre_max), int(post_max), peaks.data_ptr(),
                                          int(peaks.shape[1]), counts.data_ptr(), _stream(act.device.index)))
    launch_count += 2 if act.shape[0] else 1


def complex_flux(lgd, diff, band_lo, band_hi, temporal_filter=3, rows_per_signal=0, seg_offsets=None):
    """mmb_complex_flux: lgd (rows, K), diff (rows, B) float32 CUDA tensors, band_lo/hi int32 CUDA tensors [B].
    Returns the (rows,) onset detection function on the device."""
    global launch_count
    import torch
    rows, K = lgd.shape
    B = diff.shape[1]
    odf = torch.empty((rows,), dtype=torch.float32, device=lgd.device)
    n_seg = 0 if seg_offsets is None else int(seg_offsets.shape[0]) - 1
    rc = load().mmb_complex_flux(lgd.data_ptr(), lgd.stride(0), diff.data_ptr(), diff.stride(0), band_lo.data_ptr(),
                                 band_hi.data_ptr(), B, K, rows, int(rows_per_signal),
                                 None if seg_offsets is None else seg_offsets.data_ptr(), n_seg,
                                 int(temporal_filter), odf.data_ptr(), _stream(lgd.device.index))
    if rc != MMB_OK:
        raise ValueError("mmb_complex_flux: invalid argument") if rc == MMB_E_INVALID else LibraryError("mmb_complex_flux failed")
    launch_count += 1 if rows else 0
    return odf


ODF_MODES = {"high_frequency_content": 1, "spectral_diff": 2, "modified_kullback_leibler": 3, "phase_deviation": 4,
             "weighted_phase_deviation": 5, "normalized_weighted_phase_deviation": 6, "complex_domain": 7,
             "rectified_complex_domain": 8}


def spectral_odf(mode, stft=None, spec=None, diff_frames=1, epsilon=2.220446049250313e-16, rows_per_signal=0,
                 seg_offsets=None):
    """mmb_spectral_odf: `mode` a key of ODF_MODES; `spec` (rows, cols) float32 or `stft` (rows, K) complex64 CUDA
    tensor (row stride arbitrary, columns contiguous).  Returns the (rows,) float32 function on the device."""
    global launch_count
    import torch
    m = ODF_MODES[mode]
    src = spec if m <= 3 else stft
    if src is None or src.dim() != 2 or src.stride(1) != 1:
        raise ValueError("%s needs a 2-D %s tensor with contiguous columns" % (mode, "spec" if m <= 3 else "stft"))
    want = torch.float32 if m <= 3 else torch.complex64
    if src.dtype != want or not src.is_cuda:
        raise ValueError("%s needs a CUDA tensor of dtype %s" % (mode, want))
    rows, cols = src.shape
    odf = torch.empty((rows,), dtype=torch.float32, device=src.device)
    n_seg = 0 if seg_offsets is None else int(seg_offsets.shape[0]) - 1
    rc = load().mmb_spectral_odf(m, src.data_ptr() if m > 3 else None, src.stride(0) if m > 3 else 0,
                                 src.data_ptr() if m <= 3 else None, src.stride(0) if m <= 3 else 0, rows, cols,
                                 int(rows_per_signal), None if seg_offsets is None else seg_offsets.data_ptr(), n_seg,
                                 int(diff_frames), float(epsilon), odf.data_ptr(), _stream(src.device.index))
    if rc != MMB_OK:
        raise ValueError("mmb_spectral_odf: invalid argument") if rc == MMB_E_INVALID else LibraryError("mmb_spectral_odf failed")
    launch_count += 1 if rows else 0
    return odf


def phase_lgd(stft, want_phase=True, want_lgd=False):
    """mmb_phase_lgd on a (rows, cols) complex64 CUDA tensor: (phase or None, lgd or None), float32 on the device."""
    global launch_count
    import torch
    if not stft.is_cuda or stft.dtype != torch.complex64 or stft.dim() != 2 or stft.stride(1) != 1:
        raise ValueError("stft must be a 2-D complex64 CUDA tensor with contiguous columns")
    rows, cols = stft.shape
    ph = torch.empty((rows, cols), dtype=torch.float32, device=stft.device) if want_phase else None
    lg = torch.empty((rows, cols), dtype=torch.float32, device=stft.device) if want_lgd else None
    _check(load().mmb_phase_lgd(stft.data_ptr(), stft.stride(0), ph.data_ptr() if want_phase else None, cols,
                                lg.data_ptr() if want_lgd else None, cols, rows, cols, _stream(stft.device.index)))
    launch_count += 1 if rows else 0
    return ph, lg


def remix_mono(pcm, channel=None):
    """mmb_remix_mono: (num_frames, num_channels) int16 / float32 CUDA tensor -> (num_frames,) mono tensor of the
    same dtype (mean over the channels with the reference's rounding, or one channel)."""
    global launch_count
    import torch
    if not pcm.is_cuda or pcm.dim() != 2 or not pcm.is_contiguous() or pcm.dtype not in (torch.int16, torch.float32):
        raise ValueError("pcm must be a contiguous (num_frames, num_channels) int16 or float32 CUDA tensor")
    n, c = pcm.shape
    out = torch.empty((n,), dtype=pcm.dtype, device=pcm.device)
    rc = load().mmb_remix_mono(pcm.data_ptr(), pcm.element_size(), n, c, -1 if channel is None else int(channel),
                               out.data_ptr(), _stream(pcm.device.index))
    _check(rc)
    launch_count += 1 if n else 0
    return out
