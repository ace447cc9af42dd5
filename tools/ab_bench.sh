#!/bin/bash
# A/B the step kernel on the GPU box: builds variants (-D flags) into gpurun_out/variants/ HERE, then run on the box:
#   tools/ab_bench.sh build "name1:-DMG_STEP_THREADS=32 -DMG_STEP_MINBLOCKS=14" "name2:..."
#   gpurun -- tools/ab_bench.sh run
set -e
cd "$(dirname "$0")/.."
mkdir -p variants
if [ "$1" = build ]; then
  shift
  rm -f variants/*.so
  for spec in "$@"; do
    name="${spec%%:*}"; flags="${spec#*:}"
    python -m minigrid_dynamicprogramming_b200.build --force $flags -o=variants/$name.so > /dev/null
    echo "built variants/$name.so ($flags)"
  done
else
  for so in variants/*.so; do
    MGB200_LIB=$PWD/$so python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-dp --no-workloads 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$so', round(d['value']/1e9,3), 'G env-steps/s', round(d['roofline']['frac'],4), d['launch_shape'])"
  done
fi
