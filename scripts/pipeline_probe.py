#!/usr/bin/env python
"""mmb_pipeline_probe: the fused kernel's per-frame work as 16 monolithic warps against 12 FFT + 12 light warps."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import _lib
import torch
lib = _lib.load()
sms = torch.cuda.get_device_properties(0).multi_processor_count
for mode, name in ((0, "16 warps, each does everything"), (1, "12 FFT warps (136 regs) + 12 light warps (24 regs)")):
    v = ctypes.c_double(0)
    rc = lib.mmb_pipeline_probe(0, mode, ctypes.byref(v))
    print("mode %d %-52s rc %d  %.1f M frames/s = %.0f SM-clk per frame at 1965 MHz" % (
        mode, name, rc, v.value / 1e6, 1965e6 * sms / max(v.value, 1)))
print("fft only: %.1f M frames/s" % (_lib.fft_only_frames_per_s(0) / 1e6))
