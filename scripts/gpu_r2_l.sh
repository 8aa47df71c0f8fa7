#!/bin/bash
# round 2, call L: traversal stack levels in shared memory (PTC_SMEM_LEVELS; the rest in a global scratch) — L1 capacity A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_checked_build.py tests/test_wide_bvh.py -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/r2l_tests.txt
PTC_SMEM_LEVELS=2 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_checked_build.py -m gpu -q -x 2>&1 | tail -4 | tee -a gpurun_out/r2l_tests.txt
{
for s in 32 8 6 5 4 3 2; do PTC_SMEM_LEVELS=$s python scripts/profile_step.py 8 fast atrium 3; done
for s in 32 6 4 3; do PTC_SMEM_LEVELS=$s python scripts/profile_step.py 1 fast terrain 4; done
for s in 32 6 4 3; do PTC_SMEM_LEVELS=$s python scripts/profile_step.py 8 fast atrium@2 3; done
for s in 32 4; do PTC_SMEM_LEVELS=$s python scripts/profile_step.py 8 fast cornell 3; done
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2l_sweep.txt
