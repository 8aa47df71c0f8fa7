#!/bin/bash
# final build on 2 GPUs: the multi-GPU tests (C ABI context with NCCL and with peer copies, torchrun harness, stripes) and the weak line on both carriers
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -s 2>&1 | grep -v "^$" | tail -15 | tee gpurun_out/f2_tests_2gpu.txt
timeout 300 python bench.py --gpus 2 --steps 6 --warmup 3 --no-cpu-baseline --no-default-leg > gpurun_out/f2_weak_sp2.json 2> gpurun_out/f2_weak_sp2.err; python -c "
import json; d=json.load(open('gpurun_out/f2_weak_sp2.json')); print('weak sp', d['n_gpus'], d['value'], d['ms_per_step'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/f2_weak_tr2.json 2> gpurun_out/f2_weak_tr2.err; python -c "
import json; d=json.load(open('gpurun_out/f2_weak_tr2.json')); print('weak torchrun', d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'])"
timeout 300 python bench.py --gpus 2 --total-spp 256 --steps 2 --warmup 1 --no-cpu-baseline --no-default-leg > gpurun_out/f2_strong_c3_sp2.json 2> gpurun_out/f2_strong_c3_sp2.err; python -c "
import json; d=json.load(open('gpurun_out/f2_strong_c3_sp2.json')); print('strong c3 sp', d['n_gpus'], d['value'], d['ms_per_step'])"
