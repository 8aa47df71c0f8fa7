timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3
timeout 600 python bench.py --bvh device --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_devbvh_v9.json 2> gpurun_out/bench_devbvh_v9.err; tail -c 400 gpurun_out/bench_devbvh_v9.json; tail -3 gpurun_out/bench_devbvh_v9.err
