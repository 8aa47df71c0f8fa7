#!/bin/bash
# parked triangle groups: 1 (default build) / 2 / 3 slots per lane, with the vote thresholds that go with them
mkdir -p gpurun_out
L=$PWD/opencl_pathtracer_b200
run() { # lib scene env...
  lib=$1; sc=$2; shift 2
  env PTC_LIB_DIR=$L/$lib "$@" timeout 300 python scripts/profile_step.py 8 fast $sc 2 2>&1 | grep "^env\|^lanes" | sed "s/^/$sc $lib /" | tee -a gpurun_out/park_ab.txt
}
run _lib atrium
run _lib_park2 atrium
run _lib_park3 atrium
run _lib_stats2 atrium
for vn in 4 6; do for vt in 2 4 5; do run _lib_park2 atrium PTC_VOTE_NODE=$vn PTC_VOTE_TRI=$vt; done; done
run _lib_park3 atrium PTC_VOTE_NODE=4 PTC_VOTE_TRI=5
run _lib_park3 atrium PTC_VOTE_NODE=6 PTC_VOTE_TRI=5
run _lib terrain
run _lib_park2 terrain
run _lib cornell
run _lib_park2 cornell
