#!/bin/bash
# run bench.py for every built variant (GPU box)
cd "$(dirname "$0")/.."
for d in scratch/variants/*/; do
  n=$(basename $d)
  ARPEGGIA_B200_LIB=$PWD/$d/libarpeggia_b200.so python bench.py --no-cpu --steps 10 2>&1 | tail -1 | python -c "
import json,sys
l=json.loads(sys.stdin.readline()); print('$n', 'step', round(l['ms_per_step'],4), 'stage', {k:round(v,4) for k,v in l['stage_ms'].items()}, 'rows', l['config']['rows'], 'pairs', l['config']['pairs_within_cutoff'])"
done
