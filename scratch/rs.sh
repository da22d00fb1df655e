#!/bin/bash
cd "$(dirname "$0")/.."
export PYTHONPATH=.
for r in 8 16 32 64 128; do
  echo "rounds $r: 1M  $(ARP_RS_ROUNDS=$r python scratch/big.py 5 5 5 2>&1 | tail -1 | python -c "import json,sys; l=json.loads(sys.stdin.readline()); print(l['frame_rows'], l['frame_stats'])")"
  echo "rounds $r: 10M $(ARP_RS_ROUNDS=$r python scratch/big.py 11 11 10 2>&1 | tail -1 | python -c "import json,sys; l=json.loads(sys.stdin.readline()); print(l['frame_rows'], l['frame_stats'])")"
done
