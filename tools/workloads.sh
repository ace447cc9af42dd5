#!/bin/bash
# step throughput of the four BASELINE workloads on one GPU (headline + configs 3/4/5 grids): gpurun -- tools/workloads.sh
cd "$(dirname "$0")/.."
for w in doorkey8 doorkey16 keycorridor mixed dynobs8; do
  python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-dp --no-workloads 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w', round(d['value']/1e9,3), 'G env-steps/s', round(d['roofline']['frac'],4), d['launch_shape'])"
done
