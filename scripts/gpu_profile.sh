# One gpurun call: launch list of the bench command + ncu --set full of k_trace (extend + connect) and k_shade, current build.
set -x
TAG=${1:-v15}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
tail -3 gpurun_out/ncu1.log
python scripts/profile_step.py 8 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 2 -c 4 -f -o gpurun_out/prof_trace_$TAG python scripts/profile_step.py 8 > gpurun_out/ncu_trace.log 2>&1
tail -n 3 gpurun_out/ncu_trace.log
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 1 -c 2 -f -o gpurun_out/prof_shade_$TAG python scripts/profile_step.py 8 > gpurun_out/ncu_shade.log 2>&1
tail -n 3 gpurun_out/ncu_shade.log
ls -la gpurun_out/*.ncu-rep
