#!/bin/bash
# generic A/B: scripts/gpu_ab.sh "<lib dirs under opencl_pathtracer_b200>" "<scenes>" [repeats]  (appends to gpurun_out/ab.txt)
mkdir -p gpurun_out
L=$PWD/opencl_pathtracer_b200
for sc in $2; do for rep in $(seq 1 ${3:-2}); do for lib in $1; do
  PTC_LIB_DIR=$L/$lib timeout 300 python scripts/profile_step.py 8 fast $sc 3 2>&1 | grep "^env\|^lanes" | sed "s/^/$sc $lib /" | tee -a gpurun_out/ab.txt
done; done; done
