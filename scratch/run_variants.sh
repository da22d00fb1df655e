#!/bin/bash
# bench every kernel-tuning variant under scratch/variants (built by scratch/variants.sh) in emit-radius mode: run_variants.sh name ...
cd "$(dirname "$0")/.."
for v in default "$@"; do
  if [ "$v" = default ]; then unset ARPEGGIA_B200_LIB; else export ARPEGGIA_B200_LIB=$PWD/scratch/variants/$v/libarpeggia_b200.so; fi
  python bench.py --no-cpu --no-counting --steps 300 --flags 1 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('$v', 'step',round(d['ms_per_step'],4),'pairs',round(d['roofline']['kernel_ms'],4),{k:round(v,4) for k,v in d['roofline']['kernel_ms_each'].items()},'grid',round(d['stage_ms']['grid'],4), d['config']['rows'])"
done
