#!/bin/bash
# quick kernel check under gpurun: the parity tests that exercise the pair stage, then a short bench.   usage: bash scratch/kq.sh [extra pytest -k]
cd "$(dirname "$0")/.."
python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 20 --warmup 5 --no-cpu --no-counting > gpurun_out/kq.json 2> gpurun_out/kq.err; tail -2 gpurun_out/kq.err
python - <<PY
import json
d=json.loads(open('gpurun_out/kq.json').read().strip().splitlines()[-1])
print('step', round(d['ms_per_step'],4), 'stage', {k:round(v,4) for k,v in d['stage_ms'].items()}, 'each', {k:round(v,4) for k,v in d['roofline']['kernel_ms_each'].items()}, 'frac', round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['ms_per_step'],3))
PY
