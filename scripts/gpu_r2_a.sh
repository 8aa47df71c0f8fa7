#!/bin/bash
# round 2, call A: tests with reordering + FAST without per-ray counters, then the reordering sweep on the bench scene
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2a_tests.txt
cat gpurun_out/r2a_tests.txt
{
P="python scripts/profile_step.py 8 fast atrium 2"
PTC_SORT_RAY=-1 PTC_SORT_HIT=-1 $P
PTC_SORT_RAY=-1 PTC_SORT_HIT=-1 PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_fmad $P
for r in 1 2 3; do PTC_SORT_RAY=$r PTC_SORT_HIT=-1 $P; done
for h in 2 3 4; do PTC_SORT_RAY=-1 PTC_SORT_HIT=$h $P; done
PTC_SORT_RAY=-1 PTC_SORT_HIT=4 PTC_SORT_MIN_BOUNCE=1 $P
PTC_SORT_RAY=3 PTC_SORT_HIT=4 $P
PTC_SORT_RAY=2 PTC_SORT_HIT=4 $P
PTC_SORT_RAY=3 PTC_SORT_HIT=4 PTC_SORT_MIN_BOUNCE=1 $P
PTC_SORT_RAY=3 PTC_SORT_HIT=4 PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_fmad $P
PTC_SORT_RAY=3 PTC_SORT_HIT=4 python scripts/profile_step.py 8 parity atrium 1
PTC_SORT_RAY=-1 PTC_SORT_HIT=-1 python scripts/profile_step.py 8 parity atrium 1
PTC_SORT_RAY=3 PTC_SORT_HIT=4 python scripts/profile_step.py 1 fast terrain 4
PTC_SORT_RAY=-1 PTC_SORT_HIT=-1 python scripts/profile_step.py 1 fast terrain 4
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2a_sweep.txt
