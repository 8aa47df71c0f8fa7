TAG=${1:-v5}
python scripts/profile_step.py 2 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 2 -c 2 -f -o gpurun_out/prof_trace_$TAG python scripts/profile_step.py 2 > gpurun_out/ncu_trace.log 2>&1
tail -n 2 gpurun_out/ncu_trace.log
