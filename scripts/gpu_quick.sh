set -x
timeout 1200 python bench.py --workload terrain --steps 3 --warmup 3 > gpurun_out/bench_c4_v8.json 2> gpurun_out/bench_c4_v8.err
tail -c 300 gpurun_out/bench_c4_v8.json; tail -5 gpurun_out/bench_c4_v8.err
timeout 600 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2_v8.json 2> gpurun_out/bench_c2_v8.err
tail -c 300 gpurun_out/bench_c2_v8.json
timeout 600 python bench.py --workload cornell --progressive 600 > gpurun_out/bench_c5_v8.json 2>&1
tail -c 600 gpurun_out/bench_c5_v8.json
