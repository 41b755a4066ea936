#!/usr/bin/env python
"""Instruction-rate probes of mmb_rate_probe: thread-level instructions per second -> warp-instructions/clk/SM."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madmom_b200 import _lib
import torch
lib = _lib.load()
sms = torch.cuda.get_device_properties(0).multi_processor_count
import pynvml
pynvml.nvmlInit()
mhz = pynvml.nvmlDeviceGetMaxClockInfo(pynvml.nvmlDeviceGetHandleByIndex(0), pynvml.NVML_CLOCK_SM)
names = {1: "FFMA 3 regs", 2: "FADD", 3: "FMUL+FADD", 4: "LDS.64 + 2 FFMA", 5: "LDS.32", 6: "SHFL", 7: "LDS+SHFL",
         8: "FADD2 (f32x2)", 9: "FFMA2 (f32x2)", 10: "FFMA2 + LDS.32 + FADD", 11: "LDS.64 (2 wavefronts)",
         12: "LDS.128 (4 wavefronts)", 13: "STS.64 + LDS.64 (4 wavefronts)", 14: "STS.64 (2 wavefronts)"}
for w in range(1, 15):
    v = ctypes.c_double(0)
    rc = lib.mmb_rate_probe(0, w, ctypes.byref(v))
    print("%2d %-24s rc %d  %.1f Ginst/s  = %.2f warp-inst/clk/SM at %d MHz" % (w, names[w], rc, v.value, v.value * 1e9 / 32 / sms / (mhz * 1e6), mhz))
