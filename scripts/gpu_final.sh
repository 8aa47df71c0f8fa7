set -x
TAG=${1:-v15}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref.err; cut -c1-200 gpurun_out/bench_ref_$TAG.json
python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; cut -c1-300 gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err
python bench.py --workload terrain --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_$TAG.json 2> gpurun_out/bench_c4.err; cut -c1-300 gpurun_out/bench_c4_$TAG.json; tail -3 gpurun_out/bench_c4.err
python bench.py --workload cornell --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2_$TAG.json 2> gpurun_out/bench_c2.err; cut -c1-300 gpurun_out/bench_c2_$TAG.json
bash scripts/gpu_profile.sh $TAG
