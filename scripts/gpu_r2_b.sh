#!/bin/bash
# round 2, call B: the config-level parity tests + kernel variants (byte -> float without the conversion pipe, 9 CTAs per SM)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -x -q -s 2>&1 | tail -40 > gpurun_out/r2b_configs.txt
cat gpurun_out/r2b_configs.txt
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_abi_host.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2b_tests.txt
{
P="python scripts/profile_step.py 8 fast atrium 3"
$P
for v in q8half ctas9 q8half9 fmad; do PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_$v $P; done
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_q8half timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_wide_bvh.py -m gpu -x -q 2>&1 | tail -3
python scripts/profile_step.py 8 fast atrium@2 3
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_q8half python scripts/profile_step.py 8 fast atrium@2 3
python scripts/profile_step.py 1 fast terrain 4
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_q8half python scripts/profile_step.py 1 fast terrain 4
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2b_sweep.txt
