#!/bin/bash
# Short GPU session: GPU suite, default bench, launch list of one emit-radius step.   usage (under gpurun): bash scratch/quick_round.sh <tag>
tag=${1:-q}
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; tail -2 gpurun_out/${tag}_tests.log
python bench.py --no-cpu > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; python - <<PY
import json
d=json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['stage_ms'], d['roofline']['kernel_ms_each'], d['e2e']['ms_per_step'], d.get('e2e_concurrent',{}).get('ms_per_job'))
PY
ARP_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --no-cpu --no-counting --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_list.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/${tag}_launches.csv')) if len(r)>5]
h=rows[0]; k=h.index('Kernel Name'); v=h.index('Metric Value')
last={}
for r in rows[1:]:
    last.setdefault(r[k].split('(')[0],[]).append(r[v])
print({n:x[-1] for n,x in last.items() if n.startswith('k_')})
PY
