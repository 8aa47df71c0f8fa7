#!/bin/bash
# round 2, call D (1 GPU): class-queue shade, checked build, failing tests of call C, mismatch diagnostic, shade occupancy A/B, ncu
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2d_tests_full.txt 2>&1; grep -v "^$" gpurun_out/r2d_tests_full.txt | grep -n "C3 \|C4 \|C1 \|FAILED\|passed\|failed\|faults\|Error" | head -40
python scripts/diag_mismatch.py cornell 512 512 3 10 > gpurun_out/r2d_diag_cornell.txt 2>&1; head -30 gpurun_out/r2d_diag_cornell.txt
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
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2d_sweep.txt
# ncu: one full capture of the three hot kernels on the bench scene (bounce 1 launches), then on the 10 M-triangle scene
python scripts/profile_step.py 8 fast atrium 1 > gpurun_out/r2d_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_shade' -s 3 -c 3 -f -o gpurun_out/r2d_ncu_atrium python scripts/profile_step.py 8 fast atrium 1 > gpurun_out/r2d_ncu_atrium.log 2>&1
tail -2 gpurun_out/r2d_ncu_atrium.log
python scripts/profile_step.py 1 fast terrain 1 > gpurun_out/r2d_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_shade' -s 0 -c 6 -f -o gpurun_out/r2d_ncu_terrain python scripts/profile_step.py 1 fast terrain 1 > gpurun_out/r2d_ncu_terrain.log 2>&1
tail -2 gpurun_out/r2d_ncu_terrain.log
ls -la gpurun_out/*.ncu-rep
