#!/bin/bash
# One GPU session: full GPU suite, default bench (+ reference arm), launch list and ncu --set full of one emit-radius step.
# usage (under gpurun): bash scratch/profile_round.sh <tag>
tag=${1:-r1q}
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; tail -2 gpurun_out/${tag}_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -1 gpurun_out/${tag}_smoke.log | cut -c1-200
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 400 gpurun_out/${tag}_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; tail -c 300 gpurun_out/${tag}_bench_ref.json
ARP_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --no-cpu --no-counting --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_list.log 2>&1
ARP_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:^k_ --launch-skip 48 --launch-count 9 -f -o gpurun_out/${tag}_full \
  python bench.py --no-cpu --no-counting --steps 2 --warmup 3 > gpurun_out/${tag}_ncu_full.log 2>&1; tail -2 gpurun_out/${tag}_ncu_full.log
