python -m pytest tests -m gpu -x -q 2>&1 | tail -3
PTC_FULL_STACK=1 python scripts/profile_step.py 4 fast atrium 2
for t in 0 1 2 3; do echo "TUNE=$t"; PTC_TUNE=$t python scripts/profile_step.py 4 fast atrium 2; done
PTC_TUNE=3 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
