#!/bin/bash
# round 2, call J (8 GPUs): multi-GPU tests on 8 devices; weak + strong scaling through the C ABI's own multi-GPU context (one process);
# C4 strong at 8; progressive frames at 8 (both carriers)
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -s 2>&1 | grep -v "^$" | tail -15 | tee gpurun_out/r2j_tests_8gpu.txt
B="timeout 300 python bench.py"
for n in 1 4 8; do $B --gpus $n --steps 6 --warmup 3 --no-cpu-baseline --no-default-leg > gpurun_out/r2j_weak_sp$n.json 2> gpurun_out/r2j_weak_sp$n.err; python -c "
import json; d=json.load(open('gpurun_out/r2j_weak_sp$n.json')); print('weak sp', d['n_gpus'], d['value'], d['ms_per_step'])"; done
for n in 1 2 4 8; do $B --gpus $n --total-spp 256 --steps 2 --warmup 1 --no-cpu-baseline --no-default-leg > gpurun_out/r2j_strong_c3_sp$n.json 2> gpurun_out/r2j_strong_c3_sp$n.err; python -c "
import json; d=json.load(open('gpurun_out/r2j_strong_c3_sp$n.json')); print('strong c3 sp', d['n_gpus'], d['value'], d['ms_per_step'])"; done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --total-spp 256 --steps 2 --warmup 1 > gpurun_out/r2j_strong_c3_tr8.json 2> gpurun_out/r2j_strong_c3_tr8.err; python -c "
import json; d=json.load(open('gpurun_out/r2j_strong_c3_tr8.json')); print('strong c3 torchrun', d['n_gpus'], d['value'], d['ms_per_step'])"
timeout 400 python bench.py --gpus 8 --workload terrain --total-spp 1024 --steps 2 --warmup 1 > gpurun_out/r2j_strong_c4_sp8.json 2> gpurun_out/r2j_strong_c4_sp8.err; python -c "
import json; d=json.load(open('gpurun_out/r2j_strong_c4_sp8.json')); print('strong c4 sp', d['n_gpus'], d['value'], d['ms_per_step'])"
timeout 200 python bench.py --workload cornell --progressive 600 --gpus 8 > gpurun_out/r2j_c5_sp8.json 2>gpurun_out/r2j_c5_sp8.err; python -c "
import json; d=json.load(open('gpurun_out/r2j_c5_sp8.json')); print('c5 sp8', d['value'], d['p99_ms'], d['interior_frames']['p50_ms'], d['interior_frames']['p99_ms'])"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --workload cornell --progressive 600 > gpurun_out/r2j_c5_tr8.json 2>gpurun_out/r2j_c5_tr8.err; python -c "
import json; d=json.load(open('gpurun_out/r2j_c5_tr8.json')); print('c5 tr8', d['value'], d['p99_ms'], d['interior_frames']['p50_ms'], d['interior_frames']['p99_ms'])"
