for cfg in "5 3" "4 3" "6 3" "5 2" "5 4" "4 2" "6 4"; do set -- $cfg; echo "VN=$1 VT=$2"; PTC_VOTE_NODE=$1 PTC_VOTE_TRI=$2 python scripts/profile_step.py 8 fast atrium 3 2>&1 | tail -1; done
for r in 2 8; do echo "REFILL=$r"; PTC_REFILL_MIN=$r python scripts/profile_step.py 8 fast atrium 3 2>&1 | tail -1; done
for c in 66 85; do echo "CARVEOUT=$c"; PTC_CARVEOUT=$c python scripts/profile_step.py 8 fast atrium 3 2>&1 | tail -1; done
