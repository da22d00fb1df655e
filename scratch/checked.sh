#!/bin/bash
# The stand-in for compute-sanitizer (closed on this pool): the parity cases of scratch/sanitize_cases.py on the CHECKED build
# (-DARP_CHECKED=1: device-side index checks, dev_common.cuh), several times over (the kernels' intra-cell order differs
# from run to run, so repeated equality with the oracle also covers the lists the lanes fill without votes).
# usage (under gpurun): bash scratch/checked.sh [repeats]      build first: bash scratch/variants.sh checked "-DARP_CHECKED=1"
cd "$(dirname "$0")/.."
n=${1:-5}
export ARPEGGIA_B200_LIB=$PWD/scratch/variants/checked/libarpeggia_b200.so
{ echo "compute-sanitizer on this pool: $(compute-sanitizer --tool memcheck true 2>&1 | tail -1 | cut -c1-200)"
  echo "checked build: nvcc -DARP_CHECKED=1 (device-side index checks 0..10, dev_common.cuh / pair_kernel.cuh / grid_kernels.cuh); library $(basename $(dirname $ARPEGGIA_B200_LIB))"
  for i in $(seq 1 $n); do
    python scratch/sanitize_cases.py 2>&1 | tail -1 | cut -c1-300
  done
  if [ -f scratch/variants/checked_selftest/libarpeggia_b200.so ]; then  # negative control: check 8 with a bound real inputs violate
    ARPEGGIA_B200_LIB=$PWD/scratch/variants/checked_selftest/libarpeggia_b200.so python -c "
import arpeggia_b200 as A
try:
    A.Context(0).contacts_table(A.parse_pdb('tests/data/1ubq.pdb'), '/', 0.1, 6.5, flags=4)
    print('negative control: NOT detected')
except A.ArpeggiaError as e:
    print('negative control (check 8 deliberately too tight) is detected:', e)
" 2>&1 | tail -1
  fi
  ARP_NO_GRAPH=1 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py tests/test_gpu_sap.py -m gpu -x -q 2>&1 | tail -2
} | tee gpurun_out/checked_build.txt
