#!/bin/bash
# bench a list of variant libraries against the in-tree one: bash scratch/kq2.sh name1 name2 ...   (scratch/variants/<name>/)
cd "$(dirname "$0")/.."
run() {
  python bench.py --steps 40 --warmup 8 --no-cpu --no-counting > gpurun_out/kq_$1.json 2> gpurun_out/kq_$1.err; tail -1 gpurun_out/kq_$1.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/kq_$1.json').read().strip().splitlines()[-1])
print('$1', 'step', round(d['ms_per_step'],4), 'stage', {k:round(v,4) for k,v in d['stage_ms'].items()}, 'each', {k:round(v,4) for k,v in d['roofline']['kernel_ms_each'].items()}, 'frac', round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['ms_per_step'],3))
PY
}
run intree
for v in "$@"; do ARPEGGIA_B200_LIB=$PWD/scratch/variants/$v/libarpeggia_b200.so run $v; done
run intree2
