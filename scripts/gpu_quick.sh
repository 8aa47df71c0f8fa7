timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for r in 1 4 8 16; do PTC_REFILL_MIN=$r timeout 300 python scripts/profile_step.py 4 fast atrium 2; done
timeout 300 python scripts/profile_step.py 4 parity atrium 2
