timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
PTC_WIDE=1 timeout 300 python scripts/profile_step.py 4 fast atrium 2
PTC_WIDE=0 timeout 300 python scripts/profile_step.py 4 fast atrium 2
