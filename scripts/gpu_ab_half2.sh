#!/bin/bash
# A/B of the packed half-precision slab test (default build) against the single-precision one (_lib_f32 = -DPTC_WIDE_HALF2=0)
mkdir -p gpurun_out
if [ "$1" = tests ]; then timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/h2_tests.txt 2>&1; tail -5 gpurun_out/h2_tests.txt; fi
F32=$PWD/opencl_pathtracer_b200/_lib_f32
for sc in atrium cornell atrium@2 terrain; do
  for rep in 1 2; do
    timeout 300 python scripts/profile_step.py 8 fast $sc 3 2>&1 | grep "^env" | sed "s/^/$sc half2 /" | tee -a gpurun_out/h2_ab.txt
    PTC_LIB_DIR=$F32 timeout 300 python scripts/profile_step.py 8 fast $sc 3 2>&1 | grep "^env" | sed "s/^/$sc f32   /" | tee -a gpurun_out/h2_ab.txt
    [ $sc = terrain ] && break
  done
done
