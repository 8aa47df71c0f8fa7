python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for r in 4 8 16 24; do PTC_REFILL_MIN=$r python scripts/profile_step.py 4 fast atrium 2; done
python scripts/profile_step.py 4 parity atrium 2
