#!/bin/bash
# round 2, call E (1 GPU): ordered class queues (k_classify); tests; shade occupancy A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -x > gpurun_out/r2e_tests_full.txt 2>&1; grep -v "^$" gpurun_out/r2e_tests_full.txt | grep -n "C3 \|C4 \|C1 \|FAILED\|passed\|failed\|faults:\|Error\|FAST (" | head -40
{
P="python scripts/profile_step.py 8 fast atrium 3"
$P
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_shade3 $P
$P
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_shade3 $P
python scripts/profile_step.py 8 parity atrium 1
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_shade3 python scripts/profile_step.py 8 parity atrium 1
python scripts/profile_step.py 8 fast atrium@2 3
python scripts/profile_step.py 1 fast terrain 4
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_shade3 python scripts/profile_step.py 1 fast terrain 4
python scripts/profile_step.py 8 fast cornell 3
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_shade3 python scripts/profile_step.py 8 fast cornell 3
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2e_sweep.txt
python bench.py --workload cornell --progressive 600 > gpurun_out/r2e_c5_n1.json 2>gpurun_out/r2e_c5_n1.err; cat gpurun_out/r2e_c5_n1.json
