#!/bin/bash
# Full single-GPU check on the box: gpurun -- bash tools/gpu_suite.sh <tag> [ab] [sanitize] [ncu]
tag=${1:-suite}; shift
out=gpurun_out/$tag; mkdir -p $out
python -m pytest tests/test_zz_no_alloc.py -x -q > $out/zz_alone.log 2>&1; echo "rc=$?" >> $out/zz_alone.log
python -m pytest tests -m gpu -q > $out/gputests.log 2>&1; echo "rc=$?" >> $out/gputests.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; echo "rc=$?" >> $out/smoke.log
python bench.py --steps 20 --warmup 5 > $out/bench.json 2> $out/bench.err; echo "rc=$?" >> $out/bench.err
python bench.py --impl reference --steps 5 --warmup 1 > $out/bench_ref.json 2>&1
python bench.py --workload dynobs8 --steps 3 --warmup 3 --no-dp --no-cpu-baseline > $out/bench_dynobs.json 2>/dev/null
for a in "$@"; do
  case $a in
    ab) bash tools/ab_bench.sh run workloads > $out/ab.txt 2>&1 ;;
    abdp) for so in minigrid_dynamicprogramming_b200/libmgb200.so variants/oldtail.so; do
            MGB200_LIB=$PWD/$so python bench.py --steps 3 --warmup 3 --no-workloads --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$so', d['dp_summary'], 'unmasked_ms', d['dp']['row_mask']['unmasked_sweep_ms'], 'solve_s', d['dp']['solve']['seconds'], 'latency', d['dp']['latency_regime']['us_per_sweep'])" >> $out/abdp.txt 2>&1
          done ;;
    bounds) MGB200_LIB=$PWD/variants/bounds.so python -m pytest tests -m gpu -q -x > $out/gputests_bounds.log 2>&1; echo "rc=$?" >> $out/gputests_bounds.log ;;
    dynprof) ncu --set full --import-source on --clock-control none -k regex:^step_many_kernel -c 3 -o $out/dyn python tools/profile_dynobs.py > $out/dyn_ncu.log 2>&1
             python tools/ncu_lines.py $out/dyn.ncu-rep minigrid_dynamicprogramming_b200/libmgb200.so step_many_kernelILb1ELi0 60 1 > $out/dyn_lines.txt 2>&1 ;;
    stepprof) ncu --set full --import-source on --clock-control none -k regex:^step_many_kernel -c 3 -o $out/stepk python tools/profile_all.py > $out/step_ncu.log 2>&1
             python tools/ncu_lines.py $out/stepk.ncu-rep minigrid_dynamicprogramming_b200/libmgb200.so step_many_kernelILb0ELi0 70 1 > $out/step_lines.txt 2>&1 ;;
    launches) ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/ncu_bench.log 2>&1 ;;
    multi*) n=${a#multi}; bash tools/gpu_multi.sh $n ${tag}_x$n ;;
    ncu) ncu --set full --clock-control none --import-source on -k regex:'^(step_many_kernel|sweep_kernel|solve_kernel)' -c 11 -o $out/all python tools/profile_all.py > $out/ncu_full.log 2>&1
         ncu -i $out/all.ncu-rep --page raw --csv > $out/all_raw.csv 2>/dev/null ;;
  esac
done
