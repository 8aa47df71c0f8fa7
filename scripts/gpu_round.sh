# One gpurun call: GPU parity tests, smoke, bench (plain), ncu launch list, ncu --set full of k_trace.
set -x
TAG=${1:-v4}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/tests_$TAG.log; tail -5 gpurun_out/tests_$TAG.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --steps 4 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err
tail -c 3500 gpurun_out/bench_$TAG.json; tail -5 gpurun_out/bench_$TAG.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2>> gpurun_out/bench_$TAG.err; tail -c 800 gpurun_out/bench_ref_$TAG.json
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
tail -3 gpurun_out/ncu1.log
python scripts/profile_step.py 8 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 2 -c 4 -f -o gpurun_out/prof_trace_$TAG python scripts/profile_step.py 8 > gpurun_out/ncu_trace.log 2>&1
tail -n 3 gpurun_out/ncu_trace.log
