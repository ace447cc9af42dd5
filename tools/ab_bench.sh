#!/bin/bash
# A/B the step kernel on the GPU box: builds variants (-D flags) into variants/ HERE, then run on the box:
#   tools/ab_bench.sh build "name1:-DMG_STEP_FORCE_LAYOUT=2 -DMG_STEP_MINBLOCKS=24" "name2:..."
#   gpurun -- tools/ab_bench.sh run [workloads]        # "workloads": also the configs 3/4/5 grids
set -e
cd "$(dirname "$0")/.."
mkdir -p variants
if [ "$1" = build ]; then
  shift
  for spec in "$@"; do
    name="${spec%%:*}"; flags="${spec#*:}"
    python -m minigrid_dynamicprogramming_b200.build --force $flags -o=variants/$name.so > /dev/null
    echo "built variants/$name.so ($flags)"
  done
else
  extra="--no-workloads"; [ "$2" = workloads ] && extra=""
  for so in minigrid_dynamicprogramming_b200/libmgb200.so variants/*.so; do
    case $so in *timing*|*bounds*) continue;; esac
    MGB200_LIB=$PWD/$so python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-dp $extra 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1])
w=d.get('workloads') or {}
print('$so', round(d['value']/1e9,3), 'G env-steps/s frac', round(d['roofline']['frac'],4), 'ctas/sm', d['launch_shape']['resident_ctas_per_sm'], 'smem', d['launch_shape']['smem_per_cta'],
      {k:(round(v['value']/1e9,2), v['resident_ctas_per_sm']) for k,v in w.items()}, 'e2e', round(d['e2e']['value']/1e6,1))"
  done
fi
