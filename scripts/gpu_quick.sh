python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python scripts/profile_step.py 4 fast atrium 2
python scripts/profile_step.py 4 parity atrium 2
