tag=s11; out=gpurun_out/$tag; mkdir -p $out
python -m pytest tests -m gpu -q -x > $out/gputests.log 2>&1; echo "rc=$?" >> $out/gputests.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; echo "rc=$?" >> $out/smoke.log
for so in minigrid_dynamicprogramming_b200/libmgb200.so variants/resetold.so; do
  MGB200_LIB=$PWD/$so python bench.py --workload dynobs8 --steps 3 --warmup 3 --no-dp --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$so dynobs8', round(d['value']/1e9,3))" >> $out/ab.txt
done
bash tools/ab_bench.sh run workloads >> $out/ab.txt 2>&1
ncu --set full --import-source on --clock-control none -k regex:^step_many_kernel -c 3 -o $out/dyn python tools/profile_dynobs.py > $out/dyn_ncu.log 2>&1
tail -n 3 $out/gputests.log $out/smoke.log; cat $out/ab.txt
