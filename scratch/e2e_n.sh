#!/bin/bash
# e2e diagnosis at N ranks: per-rank host wall clock of the phases of one arp_contacts call, with / without core binding.  usage: bash scratch/e2e_n.sh N
n=${1:-4}
cd "$(dirname "$0")/.."
for mode in bind nobind; do
  [ $mode = bind ] && export ARP_BENCH_BIND=1 || unset ARP_BENCH_BIND
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 20 --warmup 5 --multi off 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']
print('$mode', 'e2e ms', round(e['ms_per_step'],3), e['cpu_binding'])
for k,v in e['last_call_per_rank'].items(): print('   ',k,v)
"
done
