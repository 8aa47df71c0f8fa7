set -x
python scripts/profile_step.py 2 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_connect -s 1 -c 2 -f -o gpurun_out/prof_connect python scripts/profile_step.py 2 > gpurun_out/ncu_connect.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_extend -s 1 -c 2 -f -o gpurun_out/prof_extend python scripts/profile_step.py 2 > gpurun_out/ncu_extend.log 2>&1
tail -3 gpurun_out/ncu_connect.log gpurun_out/ncu_extend.log
ls -la gpurun_out
