timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 300 python scripts/profile_step.py 8 fast atrium 2 | grep Msam
timeout 300 python scripts/profile_step.py 8 fast cornell 2 | grep Msam
