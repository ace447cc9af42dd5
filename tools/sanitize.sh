#!/bin/bash
# compute-sanitizer pass over the hot kernels (gpurun -- bash tools/sanitize.sh); summaries -> gpurun_out/sanitizer/
cd "$(dirname "$0")/.."
out=gpurun_out/sanitizer; mkdir -p $out
CS=/usr/local/cuda/bin/compute-sanitizer
for mode in step solve enum render wrappers; do
  for tool in memcheck racecheck; do
    timeout 900 $CS --tool $tool --print-limit 20 python tools/sanitize_targets.py $mode > $out/${tool}_$mode.log 2>&1
    echo "$tool $mode rc=$? :: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $out/${tool}_$mode.log | tail -1)" >> $out/summary.txt
  done
done
cat $out/summary.txt
