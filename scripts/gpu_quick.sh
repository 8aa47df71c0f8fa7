timeout 300 python bench.py --workload cornell --progressive 600 > gpurun_out/bench_c5_v11.json 2> gpurun_out/bench_c5_v11.err; tail -c 1000 gpurun_out/bench_c5_v11.json | cut -c1-400; tail -3 gpurun_out/bench_c5_v11.err
PTC_GRAPHS=0 timeout 300 python bench.py --workload cornell --progressive 600 2>/dev/null | cut -c1-400
