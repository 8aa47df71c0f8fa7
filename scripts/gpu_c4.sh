TAG=${1:-v12}
timeout 1500 python bench.py --workload terrain --steps 3 --warmup 3 > gpurun_out/bench_c4_$TAG.json 2> gpurun_out/bench_c4_$TAG.err
cut -c1-600 gpurun_out/bench_c4_$TAG.json; tail -3 gpurun_out/bench_c4_$TAG.err
timeout 600 python bench.py --workload cornell --steps 4 --warmup 3 > gpurun_out/bench_c2_$TAG.json 2> gpurun_out/bench_c2_$TAG.err
cut -c1-400 gpurun_out/bench_c2_$TAG.json; tail -3 gpurun_out/bench_c2_$TAG.err
timeout 900 python bench.py --workload terrain --bvh device --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_devbvh_$TAG.json 2> gpurun_out/bench_c4_devbvh_$TAG.err
cut -c1-400 gpurun_out/bench_c4_devbvh_$TAG.json; tail -3 gpurun_out/bench_c4_devbvh_$TAG.err
