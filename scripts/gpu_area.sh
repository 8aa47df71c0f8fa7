set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v13b.json 2> gpurun_out/bench_v13b.err; python -c "
import json; d=json.load(open('gpurun_out/bench_v13b.json')); print(d['value'], d['stage_ms_per_step'])"
python bench.py --workload cornell --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2_v13.json 2> gpurun_out/bench_c2.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c2_v13.json')); print(d['value'], d['mrays_per_s'], d['stage_ms_per_step'])"
python bench.py --workload cornell_area --steps 4 --warmup 3 > gpurun_out/bench_c2_area_v13.json 2> gpurun_out/bench_c2a.err; tail -3 gpurun_out/bench_c2a.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c2_area_v13.json')); print(d['value'], d['mrays_per_s'], d['stage_ms_per_step'], d['mean_rgb'])"
