#!/bin/bash
# Round-2 evidence in one GPU session: launch list, the same restricted to the NVTX range of the pair stage, ncu --set full of
# every kernel of one emit-radius step.     usage (under gpurun): bash scratch/profile_r2.sh <tag>
tag=${1:-r2b}
cd "$(dirname "$0")/.."
CMD="python bench.py --no-cpu --no-counting --steps 2 --warmup 3"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { echo "plain run failed"; tail -3 gpurun_out/${tag}_plain.err; exit 1; }
ARP_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_list.log 2>&1
ARP_NO_GRAPH=1 ncu --nvtx --nvtx-include "arp:pairs+rings/" --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${tag}_launches_nvtx_pairs.csv $CMD > gpurun_out/${tag}_ncu_nvtx.log 2>&1
ARP_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:^k_ --launch-skip 48 --launch-count 9 -f -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu_full.log 2>&1; tail -2 gpurun_out/${tag}_ncu_full.log
python - <<PY
import csv
for f in ('gpurun_out/${tag}_launches.csv', 'gpurun_out/${tag}_launches_nvtx_pairs.csv'):
    rows=[r for r in csv.reader(open(f)) if len(r)>5]
    h=rows[0]; k=h.index('Kernel Name'); v=h.index('Metric Value')
    names=[r[k].split('(')[0] for r in rows[1:]]
    print(f, len(names), 'launches; last 12:', list(zip(names[-12:], [r[v] for r in rows[-12:]])))
PY
