#!/bin/bash
# round 2, call F: where does the class-queue shade spend its time? launch list + one full capture of k_classify / k_shade
mkdir -p gpurun_out
python scripts/profile_step.py 8 fast atrium 1 > gpurun_out/r2f_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2f_launches.csv python scripts/profile_step.py 8 fast atrium 1 > gpurun_out/r2f_ncu1.log 2>&1
python scripts/summarize_launches.py gpurun_out/r2f_launches.csv | tee gpurun_out/r2f_launches_summary.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_classify|k_shade' -s 2 -c 4 -f -o gpurun_out/r2f_ncu_shade python scripts/profile_step.py 8 fast atrium 1 > gpurun_out/r2f_ncu2.log 2>&1
tail -2 gpurun_out/r2f_ncu2.log
