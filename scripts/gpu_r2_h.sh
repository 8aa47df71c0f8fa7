#!/bin/bash
# round 2, call G: tile-partitioned queue; tests; shade register A/B; launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2h_tests.txt 2>&1; tail -4 gpurun_out/r2h_tests.txt
{
for v in "" shade3; do
  export PTC_LIB_DIR=${v:+$PWD/opencl_pathtracer_b200/_lib_$v}; [ -z "$v" ] && unset PTC_LIB_DIR
  python scripts/profile_step.py 8 fast atrium 3
  python scripts/profile_step.py 8 fast cornell 3
  python scripts/profile_step.py 1 fast terrain 4
done
unset PTC_LIB_DIR
python scripts/profile_step.py 8 parity atrium 1
python scripts/profile_step.py 8 fast atrium@2 3
} 2>&1 | grep -v "^scene" | tee gpurun_out/r2h_sweep.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 140 --csv --log-file gpurun_out/r2h_launches.csv python scripts/profile_step.py 8 fast atrium 1 > gpurun_out/r2h_ncu1.log 2>&1
python scripts/summarize_launches.py gpurun_out/r2h_launches.csv | tee gpurun_out/r2h_launches_summary.txt
python bench.py --workload cornell --progressive 600 > gpurun_out/r2h_c5_n1.json 2>gpurun_out/r2h_c5_n1.err; cat gpurun_out/r2h_c5_n1.json | head -c 600
