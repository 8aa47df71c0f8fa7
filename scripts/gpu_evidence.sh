#!/bin/bash
# evidence for the final build: (1) the whole GPU suite on the CHECKED library (bounds checks on every index; does any parity test trip one?),
# (2) ncu of the PARITY kernels (the bit-exact mode), (3) ncu of the 1 M-triangle atrium
mkdir -p gpurun_out
PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_checked timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_checked_build.py > gpurun_out/ev_checked_suite.txt 2>&1; tail -3 gpurun_out/ev_checked_suite.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_shade' -s 3 -c 3 -f -o gpurun_out/ev_ncu_parity python scripts/profile_step.py 2 parity atrium 1 > gpurun_out/ev_ncu_parity.log 2>&1; tail -1 gpurun_out/ev_ncu_parity.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_shade' -s 3 -c 3 -f -o gpurun_out/ev_ncu_atrium1m python scripts/profile_step.py 8 fast atrium@2 1 > gpurun_out/ev_ncu_atrium1m.log 2>&1; tail -1 gpurun_out/ev_ncu_atrium1m.log
