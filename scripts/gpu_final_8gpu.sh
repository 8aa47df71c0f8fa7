#!/bin/bash
# final build on 8 GPUs: weak (both carriers) and strong (C3 256 spp, C4 1024 spp at 4K) lines
mkdir -p gpurun_out
nvidia-smi -L | wc -l
B="timeout 300 python bench.py"
$B --gpus 8 --steps 6 --warmup 3 --no-cpu-baseline --no-default-leg > gpurun_out/f8_weak_sp8.json 2> gpurun_out/f8_weak_sp8.err; python -c "
import json; d=json.load(open('gpurun_out/f8_weak_sp8.json')); print('weak sp', d['n_gpus'], d['value'], d['ms_per_step'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/f8_weak_tr8.json 2> gpurun_out/f8_weak_tr8.err; python -c "
import json; d=json.load(open('gpurun_out/f8_weak_tr8.json')); print('weak torchrun', d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'])"
$B --gpus 8 --total-spp 256 --steps 2 --warmup 1 --no-cpu-baseline --no-default-leg > gpurun_out/f8_strong_c3_sp8.json 2> gpurun_out/f8_strong_c3_sp8.err; python -c "
import json; d=json.load(open('gpurun_out/f8_strong_c3_sp8.json')); print('strong c3 sp', d['n_gpus'], d['value'], d['ms_per_step'])"
timeout 400 python bench.py --gpus 8 --workload terrain --total-spp 1024 --steps 2 --warmup 1 --no-cpu-baseline --no-default-leg > gpurun_out/f8_strong_c4_sp8.json 2> gpurun_out/f8_strong_c4_sp8.err; python -c "
import json; d=json.load(open('gpurun_out/f8_strong_c4_sp8.json')); print('strong c4 sp', d['n_gpus'], d['value'], d['ms_per_step'], d.get('mrays_per_s'))"
