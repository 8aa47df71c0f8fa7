#!/bin/bash
# round 2, call K: pooled triangle tests (PTC_TRIQUEUE=1) — correctness, then A/B
mkdir -p gpurun_out
PTC_TRIQUEUE=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_checked_build.py -m gpu -q -x 2>&1 | tail -15 | tee gpurun_out/r2k_tests_tq.txt
{
python scripts/profile_step.py 8 fast atrium 3
PTC_TRIQUEUE=1 python scripts/profile_step.py 8 fast atrium 3
for vn in 3 4 6 7; do PTC_TRIQUEUE=1 PTC_VOTE_NODE=$vn python scripts/profile_step.py 8 fast atrium 3; done
PTC_TRIQUEUE=1 python scripts/profile_step.py 8 fast cornell 3
python scripts/profile_step.py 8 fast cornell 3
PTC_TRIQUEUE=1 python scripts/profile_step.py 1 fast terrain 4
PTC_TRIQUEUE=1 python scripts/profile_step.py 8 fast atrium@2 3
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2k_sweep.txt
