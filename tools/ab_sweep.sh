#!/bin/bash
# A/B the Bellman sweep on the GPU box: build variants HERE with tools/ab_bench.sh build "name:-Dflags" ..., then
#   gpurun -- tools/ab_sweep.sh
cd "$(dirname "$0")/.."
for so in variants/*.so; do
  MGB200_LIB=$PWD/$so python tools/sweep_bench.py 2>&1 | tail -2
done
