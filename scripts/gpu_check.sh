#!/bin/bash
# One GPU-box call: GPU test-suite, default bench line, FFT-only probe.  Outputs under gpurun_out/<tag>_*.
tag=${1:-run}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/${tag}_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench rc=$?"
python -c "
from madmom_b200 import _lib
print('fft_only frames/s', _lib.fft_only_frames_per_s(0), 'fp32 peak', _lib.measure_fp32_peak(0))
" > gpurun_out/${tag}_probe.txt 2>&1
cat gpurun_out/${tag}_probe.txt
python - <<'PY'
import json,sys
try:
    l=json.loads(open('gpurun_out/'+sys.argv[1]+'_bench.json').read().strip().splitlines()[-1]) if False else None
except Exception as e:
    pass
PY
tail -c 3000 gpurun_out/${tag}_bench.json
