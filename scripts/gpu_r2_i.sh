#!/bin/bash
# round 2, call I: per-scene partition switch; tests (both settings); shade register A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2i_tests.txt 2>&1; tail -3 gpurun_out/r2i_tests.txt
PTC_PARTITION=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -2 | tee gpurun_out/r2i_tests_part1.txt
PTC_PARTITION=0 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -2 | tee gpurun_out/r2i_tests_part0.txt
{
for v in "" shade3; do
  export PTC_LIB_DIR=${v:+$PWD/opencl_pathtracer_b200/_lib_$v}; [ -z "$v" ] && unset PTC_LIB_DIR
  python scripts/profile_step.py 8 fast atrium 3
  python scripts/profile_step.py 8 fast cornell 3
  python scripts/profile_step.py 1 fast terrain 4
  PTC_PARTITION=1 python scripts/profile_step.py 1 fast terrain 4
  python scripts/profile_step.py 8 fast atrium@2 3
done
unset PTC_LIB_DIR
python scripts/profile_step.py 8 parity atrium 1
python scripts/profile_step.py 8 parity cornell 1
PTC_PARTITION=0 python scripts/profile_step.py 8 parity cornell 1
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2i_sweep.txt
