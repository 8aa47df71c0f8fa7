#!/bin/bash
# round 2, call C (2 GPUs): all tests incl. the multi-GPU ones, sanitizer runs, connect occupancy A/B, first bench lines
mkdir -p gpurun_out
nvidia-smi -L | tee gpurun_out/r2c_gpus.txt
timeout 1500 python -m pytest tests -m gpu -q -s 2>&1 | tail -60 > gpurun_out/r2c_tests.txt
tail -30 gpurun_out/r2c_tests.txt
{
P="python scripts/profile_step.py 8 fast atrium 3"
$P
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_any7 $P
$P
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_any7 $P
python scripts/profile_step.py 8 parity atrium 1
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2c_sweep.txt
python bench.py --steps 10 --warmup 3 > gpurun_out/r2c_bench_n1.json 2> gpurun_out/r2c_bench_n1.err; tail -c 1500 gpurun_out/r2c_bench_n1.json
python bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2c_bench_sp2.json 2> gpurun_out/r2c_bench_sp2.err; tail -c 800 gpurun_out/r2c_bench_sp2.json; tail -3 gpurun_out/r2c_bench_sp2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2c_bench_tr2.json 2> gpurun_out/r2c_bench_tr2.err; tail -c 600 gpurun_out/r2c_bench_tr2.json
python bench.py --gpus 2 --total-spp 64 --steps 3 --warmup 1 > gpurun_out/r2c_bench_sp2_strong.json 2> gpurun_out/r2c_bench_sp2_strong.err; tail -c 600 gpurun_out/r2c_bench_sp2_strong.json
python bench.py --gpus 1 --total-spp 64 --steps 3 --warmup 1 --no-cpu-baseline --no-default-leg > gpurun_out/r2c_bench_n1_strong.json 2> gpurun_out/r2c_bench_n1_strong.err; tail -c 400 gpurun_out/r2c_bench_n1_strong.json
python bench.py --workload cornell --progressive 600 > gpurun_out/r2c_c5_n1.json 2>gpurun_out/r2c_c5_n1.err; cat gpurun_out/r2c_c5_n1.json
python bench.py --workload cornell --progressive 600 --gpus 2 > gpurun_out/r2c_c5_sp2.json 2>gpurun_out/r2c_c5_sp2.err; cat gpurun_out/r2c_c5_sp2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --workload cornell --progressive 600 > gpurun_out/r2c_c5_tr2.json 2>gpurun_out/r2c_c5_tr2.err; cat gpurun_out/r2c_c5_tr2.json
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_case.py > gpurun_out/r2c_sanitizer_$tool.txt 2>&1
  tail -8 gpurun_out/r2c_sanitizer_$tool.txt
done
