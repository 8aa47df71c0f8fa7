set -x
timeout 1200 python bench.py --workload terrain --steps 3 --warmup 3 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
tail -c 3000 gpurun_out/bench_c4.json; tail -5 gpurun_out/bench_c4.err
