for ex in 1 0 -1 1 0 -1; do PTC_STACK_EXTRA=$ex timeout 300 python scripts/profile_step.py 8 fast atrium 3 2>&1 | grep "^env" | tee -a gpurun_out/stack.txt; done
for ex in 1 0; do PTC_STACK_EXTRA=$ex timeout 300 python scripts/profile_step.py 8 fast terrain 3 2>&1 | grep "^env" | tee -a gpurun_out/stack.txt; done
for ex in 1 0; do PTC_STACK_EXTRA=$ex timeout 300 python scripts/profile_step.py 8 fast atrium@2 3 2>&1 | grep "^env" | tee -a gpurun_out/stack.txt; done
PTC_STACK_EXTRA=-1 PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_checked timeout 300 python -m pytest tests/test_checked_build.py -m gpu -q 2>&1 | tail -3
