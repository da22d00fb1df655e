#!/bin/bash
# compute-sanitizer over scratch/sanitize_cases.py, one tool per run; summaries -> gpurun_out/sanitizer_<tool>.txt     usage (under gpurun): bash scratch/sanitize.sh [tools...]
cd "$(dirname "$0")/.."
python scratch/sanitize_cases.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "cases fail WITHOUT the sanitizer:"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
tail -1 gpurun_out/sanitize_plain.log
for tool in ${@:-memcheck racecheck initcheck synccheck}; do
  extra=""
  [ "$tool" = initcheck ] && extra="--track-unused-memory no"
  start=$(date +%s)
  timeout 1500 compute-sanitizer --tool $tool $extra --print-limit 40 --log-file gpurun_out/sanitizer_${tool}_full.log python scratch/sanitize_cases.py > gpurun_out/sanitizer_${tool}_stdout.log 2>&1
  rc=$?
  { echo "compute-sanitizer --tool $tool $extra python scratch/sanitize_cases.py   (exit code $rc, $(( $(date +%s) - start )) s)"
    tail -1 gpurun_out/sanitizer_${tool}_stdout.log | cut -c1-400
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|Uninitialized|Error:" gpurun_out/sanitizer_${tool}_full.log | sort | uniq -c | sort -rn | head -30
  } > gpurun_out/sanitizer_${tool}.txt
  cat gpurun_out/sanitizer_${tool}.txt | head -12
done
