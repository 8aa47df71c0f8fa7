timeout 900 python -m pytest tests -m gpu -q -x -s -k "russian" 2>&1 | grep -v "^$" | tail -12 | cut -c1-300
timeout 300 python scripts/profile_step.py 8 fast atrium 2
