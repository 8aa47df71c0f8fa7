for s in 1 2 4 8 16; do timeout 300 python scripts/profile_step.py $s fast atrium 3; done
