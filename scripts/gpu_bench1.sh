set -x
python -c "import __graft_entry__ as g; g.smoke()"
python bench.py --steps 4 --warmup 3 > gpurun_out/bench1.json 2> gpurun_out/bench1.err
tail -c 3000 gpurun_out/bench1.json; tail -5 gpurun_out/bench1.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
tail -3 gpurun_out/ncu1.log
