"""Quick device-side throughput probe (not the bench): frames/s of the fused kernel for a few recipes."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from madmom_b200 import _lib
from madmom_b200.synth import make_batch_torch
from oracle import frontend as O

def probe(N, bpo, norm, max_bins, n_tracks=256, seconds=30.0, hop=441., outs=("diff", "odf"), reps=5, dtype=np.int16):
    win = np.hanning(N)
    fb = O.logarithmic_filterbank(O.fft_frequencies(N // 2, 44100), bpo, norm_filters=norm)
    d = O.diff_frames(0.5, hop, N, win)
    plan = _lib.FrontEndPlan(N, hop, O.fft_window(win, N, dtype), pcm_dtype=dtype, filterbank=fb, log_mode=_lib.LOG_LOG10,
                             mul=1.0, add=1.0, diff_frames=d, diff_max_bins=max_bins, positive_diffs=True)
    pcm = make_batch_torch(n_tracks, seconds, dtype=None if dtype == np.int16 else torch.float32)
    n = pcm.shape[1]
    T = _lib.num_frames(n, hop)
    units = _lib.make_units([n] * n_tracks, [T] * n_tracks)
    rows = T * n_tracks
    o = {}
    if "diff" in outs: o["diff"] = torch.empty((rows, plan.num_bands), device="cuda")
    if "odf" in outs: o["odf"] = torch.empty((rows,), device="cuda")
    if "log" in outs: o["log"] = torch.empty((rows, plan.num_bands), device="cuda")
    if "spec" in outs: o["spec"] = torch.empty((rows, plan.num_bins), device="cuda")
    pcm = pcm.reshape(-1)
    for _ in range(2): plan.run(pcm, units, o)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): plan.run(pcm, units, o)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    fps = rows / (ms * 1e-3)
    print("N=%d bpo=%d bands=%d outs=%s dtype=%s: %.3f ms/launch, %.1f Mframes/s, %.2f M x realtime, info=%s"
          % (N, bpo, plan.num_bands, ",".join(outs), np.dtype(dtype).name, ms, fps / 1e6, fps * hop / 44100 / 1e6, plan.info), flush=True)

if __name__ == "__main__":
    print("fp32 peak TFLOP/s:", _lib.measure_fp32_peak(0), flush=True)
    probe(2048, 24, False, 3)
    probe(2048, 12, False, 0)
    probe(2048, 12, False, 0, outs=("odf",))
    probe(2048, 12, False, 0, outs=("spec",), n_tracks=64)
    probe(1024, 6, True, 0, outs=("log", "diff"))
    probe(4096, 6, True, 0, outs=("log", "diff"))
    probe(2048, 24, False, 3, dtype=np.float32, n_tracks=128)
