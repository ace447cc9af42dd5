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
    sanitize) bash tools/sanitize.sh > $out/sanitize.txt 2>&1; cp -r gpurun_out/sanitizer $out/ 2>/dev/null ;;
    ncu) ncu --set full --clock-control none --import-source on -k regex:'^(step_many_kernel|sweep_kernel|solve_kernel)' -c 9 -o $out/all python tools/profile_all.py > $out/ncu_full.log 2>&1
         ncu -i $out/all.ncu-rep --page raw --csv > $out/all_raw.csv 2>/dev/null ;;
  esac
done
