mkdir -p gpurun_out/r2a
nvidia-smi topo -m > gpurun_out/r2a/topo.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/r2a/gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a/gputests.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2a/bench.json 2> gpurun_out/r2a/bench.err; echo "bench rc=$?" >> gpurun_out/r2a/bench.err
for c in 1 2 4 8; do MG_HOST_CHUNKS=$c python bench.py --steps 10 --warmup 3 --no-dp --no-workloads --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunks $c e2e', round(d['e2e']['value']/1e6,1), 'M env-steps/s', d['e2e']['d2h_gbs'], d['e2e']['d2h_ceiling_gbs'], 'value', round(d['value']/1e9,3))" >> gpurun_out/r2a/e2e_chunks.txt; done
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2a/bench_ref.json 2>&1
