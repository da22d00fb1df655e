#!/bin/bash
# bench every kernel-tuning variant under scratch/variants (built by scratch/variants.sh) in emit-radius mode
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q -k "emit_radius or screen_modes or edge_cases" > gpurun_out/t5.log 2>&1; tail -2 gpurun_out/t5.log
for v in default "$@"; do
  if [ "$v" = default ]; then unset ARPEGGIA_B200_LIB; else export ARPEGGIA_B200_LIB=$PWD/scratch/variants/$v/libarpeggia_b200.so; fi
  python bench.py --no-cpu --no-counting --steps 200 --flags 1 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('$v', 'step',round(d['ms_per_step'],4),'pairs',round(d['roofline']['kernel_ms'],4),{k:round(v,4) for k,v in d['roofline']['kernel_ms_each'].items()},'grid',round(d['stage_ms']['grid'],4),'e2e',round(d['e2e']['ms_per_step'],3),'q',d['config']['pair_queue_entries'], d['config']['rows'])"
done
