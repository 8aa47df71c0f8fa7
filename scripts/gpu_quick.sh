timeout 300 python scripts/profile_step.py 8 fast atrium 2 | grep Msam
