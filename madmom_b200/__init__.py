"""madmom_b200 -- a B200-native (sm_100a) spectral front end behind madmom's Processor API.

The package holds only what the hot path needs: `csrc/` (CUDA kernels + the C-ABI library
`libmadmom_b200.so`), the ctypes binding (`_lib`) and the host-side mirror of the reference's
processor/result classes (`processors`, `audio.*`, `features.onsets`).  There is no CPU fallback: every
numeric stage runs in the CUDA library and importing a compute path without it fails loudly.
"""
__version__ = "0.1.0"
