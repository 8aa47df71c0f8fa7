python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for r in 1 4 8 12 16 24 32; do PTC_REFILL_MIN=$r python scripts/profile_step.py 4 fast atrium 2; done
for c in 4 6 8; do PTC_CTAS_PER_SM=$c python scripts/profile_step.py 4 fast atrium 2; done
python scripts/profile_step.py 4 parity atrium 2
