#!/bin/bash
# Build kernel-tuning variants of libarpeggia_b200.so: scratch/variants.sh name "EXTRA flags" ...
set -e
cd "$(dirname "$0")/.."
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  out=$PWD/scratch/variants/$name
  mkdir -p $out
  make -s -C arpeggia_b200/csrc -j8 OUT=$out EXTRA="$flags" 2>&1 | grep -i "error" || true
  grep -A2 "k_screen_atoms" $out/ptxas.log | tail -2 | tr '\n' ' '; echo " <- $name"
done
