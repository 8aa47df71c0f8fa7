#!/bin/bash
mkdir -p gpurun_out
for cfg in "5 3 4" "4 3 4" "6 3 4" "5 2 4" "5 4 4" "5 3 2" "5 3 6" "5 3 8"; do
  set -- $cfg
  PTC_VOTE_NODE=$1 PTC_VOTE_TRI=$2 PTC_REFILL_MIN=$3 timeout 300 python scripts/profile_step.py 8 fast atrium 3 2>&1 | grep "^env" | tee -a gpurun_out/votes.txt
done
