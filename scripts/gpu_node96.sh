set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for i in 1 2; do python scripts/profile_step.py 8 fast atrium 3 2>&1 | tail -1; done
python scripts/profile_step.py 1 fast terrain 3 2>&1 | tail -1
python scripts/profile_step.py 8 fast cornell 3 2>&1 | tail -1
