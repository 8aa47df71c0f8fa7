#!/bin/bash
# round 2, final measurement call (1 GPU): the whole GPU suite, smoke, the bench lines of every named configuration, the reference arm,
# the ncu launch list of the default bench command and one full capture of the hot kernels (atrium, Cornell)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/m1_tests_full.txt 2>&1; grep -v "^$" gpurun_out/m1_tests_full.txt | grep "C3 \|C4 \|C1 \|FAILED\|passed\|failed\|faults:\|FAST (" | head -40
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/m1_smoke.txt
B="timeout 600 python bench.py"
$B --steps 10 --warmup 3 > gpurun_out/m1_bench_atrium.json 2> gpurun_out/m1_bench_atrium.err; tail -c 300 gpurun_out/m1_bench_atrium.json; echo
$B --impl reference --steps 4 --warmup 1 > gpurun_out/m1_bench_ref_atrium.json 2> gpurun_out/m1_bench_ref_atrium.err
$B --workload atrium_1m --steps 8 --warmup 3 > gpurun_out/m1_bench_atrium_1m.json 2> gpurun_out/m1_bench_atrium_1m.err
$B --workload cornell --steps 10 --warmup 3 > gpurun_out/m1_bench_cornell.json 2> gpurun_out/m1_bench_cornell.err
$B --workload cornell512 --steps 10 --warmup 3 > gpurun_out/m1_bench_cornell512.json 2> gpurun_out/m1_bench_cornell512.err
$B --impl reference --workload cornell512 --ref-rows 512 --steps 16 --warmup 1 > gpurun_out/m1_bench_ref_cornell512.json 2> gpurun_out/m1_bench_ref_cornell512.err
$B --workload cornell_area --steps 10 --warmup 3 > gpurun_out/m1_bench_cornell_area.json 2> gpurun_out/m1_bench_cornell_area.err
$B --workload terrain --steps 20 --warmup 3 > gpurun_out/m1_bench_terrain.json 2> gpurun_out/m1_bench_terrain.err
$B --workload cornell --progressive 600 > gpurun_out/m1_bench_c5.json 2> gpurun_out/m1_bench_c5.err
for f in atrium ref_atrium atrium_1m cornell cornell512 ref_cornell512 cornell_area terrain c5; do python - <<PY
import json
try:
    d = json.load(open("gpurun_out/m1_bench_$f.json"))
    print("$f", d.get("value"), d.get("e2e", {}).get("value"), (d.get("roofline") or {}).get("frac"), (d.get("cpu_baseline") or {}).get("value"))
except Exception as e:
    print("$f ERR", e, open("gpurun_out/m1_bench_$f.err").read()[-600:])
PY
done
# ncu: the launch list of the default bench command (shares), then one full capture of extend / connect / shade of bounce 1
$B --steps 2 --warmup 1 --no-cpu-baseline --no-default-leg > gpurun_out/m1_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 400 --csv --log-file gpurun_out/m1_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-default-leg > gpurun_out/m1_ncu_launches.log 2>&1
python scripts/summarize_launches.py gpurun_out/m1_launches.csv | tee gpurun_out/m1_launches_summary.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_shade' -s 3 -c 3 -f -o gpurun_out/m1_ncu_atrium python scripts/profile_step.py 8 fast atrium 1 > gpurun_out/m1_ncu_atrium.log 2>&1; tail -1 gpurun_out/m1_ncu_atrium.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_shade|k_classify' -s 4 -c 4 -f -o gpurun_out/m1_ncu_cornell python scripts/profile_step.py 8 fast cornell 1 > gpurun_out/m1_ncu_cornell.log 2>&1; tail -1 gpurun_out/m1_ncu_cornell.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_shade' -s 0 -c 3 -f -o gpurun_out/m1_ncu_terrain python scripts/profile_step.py 1 fast terrain 1 > gpurun_out/m1_ncu_terrain.log 2>&1; tail -1 gpurun_out/m1_ncu_terrain.log
