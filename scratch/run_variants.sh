#!/bin/bash
# run bench.py for every built variant (GPU box)
cd "$(dirname "$0")/.."
for d in scratch/variants/*/; do
  n=$(basename $d)
  ARPEGGIA_B200_LIB=$PWD/$d/libarpeggia_b200.so python bench.py --no-cpu --steps 30 2>/dev/null | tail -1 > /tmp/v.json
  python - "$n" <<PY
import json,sys
l=json.loads(open('/tmp/v.json').read()); print(sys.argv[1], 'step', round(l['ms_per_step'],4), {k:round(v,4) for k,v in l['stage_ms'].items()}, {k:round(v,4) for k,v in l['roofline']['kernel_ms_each'].items()}, 'rows', l['config']['rows'], 'frame_sort', round(l['frame']['ms_sort'],3))
PY
done
