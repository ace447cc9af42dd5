#!/bin/bash
# Multi-GPU DP check + bench on the box: gpurun --gpus N -- bash tools/gpu_multi.sh N <tag>
#   third argument "benchonly": just the bench line (the DP checks / timing tools are unchanged since their last run)
N=${1:-2}; tag=${2:-multi$N}; mode=${3:-all}
mkdir -p gpurun_out/$tag
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/$tag/topo.txt 2>&1
if [ "$mode" != benchonly ]; then
$TR --master-port 29611 tools/dp_sharded_check.py > gpurun_out/$tag/sharded_check.log 2>&1; echo "rc=$?" >> gpurun_out/$tag/sharded_check.log
$TR --master-port 29617 tools/dp_push_timing.py > gpurun_out/$tag/push_new.log 2>&1; echo "rc=$?" >> gpurun_out/$tag/push_new.log
if [ -f variants/timing.so ]; then
  MGB200_LIB=$PWD/variants/timing.so MGDP_BENCH_QUICK=1 $TR --master-port 29618 tools/dp_push_timing.py > gpurun_out/$tag/push_timing_new.log 2>&1
  MGB200_LIB=$PWD/variants/timing_oldsync.so MGDP_BENCH_QUICK=1 $TR --master-port 29619 tools/dp_push_timing.py > gpurun_out/$tag/push_timing_old.log 2>&1
fi
for v in variants/push*.so variants/oldtail.so; do   # A/B builds of the push kernel (e.g. -DMGDP_SOLVE_PUSH_MINB=8)
  [ -f "$v" ] && MGB200_LIB=$PWD/$v MGDP_BENCH_QUICK=1 $TR --master-port 29622 tools/dp_push_timing.py > gpurun_out/$tag/push_$(basename $v .so).log 2>&1
done
if [ "$N" = 8 ]; then      # the same solve on 4 of the 8 GPUs (strong-scaling table)
  CUDA_VISIBLE_DEVICES=0,1,2,3 MGDP_BENCH_QUICK=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29621 tools/dp_push_timing.py > gpurun_out/$tag/push_new_4of8.log 2>&1
fi
fi
NCCL_DEBUG=INFO $TR --master-port 29620 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/$tag/bench.out 2> gpurun_out/$tag/bench.err; echo "rc=$?" >> gpurun_out/$tag/bench.err
grep '^{"metric"' gpurun_out/$tag/bench.out | tail -1 > gpurun_out/$tag/bench.json; tail -1 gpurun_out/$tag/bench.out | cut -c1-80 > gpurun_out/$tag/bench_lastline.txt
[ "$mode" != benchonly ] && python bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/$tag/bench_ref.json 2>&1
true
