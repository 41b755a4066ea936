#!/bin/bash
# Times the default workload (config 2) for each value of an environment knob:  gpu_sweep.sh VAR v1 v2 ...
# ("-" = unset).  Prints the plan info (warps, shared memory) beside the time.
var=$1; shift
for v in "$@"; do
  if [ "$v" = "-" ]; then unset $var; else export $var=$v; fi
  python - <<PY
import json, subprocess, sys, os
import numpy as np
from madmom_b200 import recipe as R
from madmom_b200.batch import BatchFrontEnd
fe = BatchFrontEnd(R.LOGFILTSPECFLUX)
info = fe.plans[0].info
out = subprocess.run([sys.executable, "bench.py", "--steps", "10", "--warmup", "3", "--no-cpu", "--no-e2e", "--no-configs"],
                     capture_output=True, text=True)
try:
    l = json.loads(out.stdout.strip().splitlines()[-1])
    print("$var=$v", "ms %.3f" % l["ms_per_step"], "frac %.4f" % l["roofline"]["frac"], "warps", info["warps"], "smem", info["smem_bytes"])
except Exception as e:
    print("$var=$v failed", out.stderr[-300:])
PY
done
