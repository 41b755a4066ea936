"""CPU oracle for the madmom spectral front end -- TEST INFRASTRUCTURE, not product code.

Only `tests/`, `__graft_entry__.smoke()` and the CPU-baseline / `--impl reference` legs of `bench.py` may
import this package.  `madmom_b200/` never imports it and has no CPU fallback.  See `oracle/frontend.py`.
"""
