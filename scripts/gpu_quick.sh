timeout 300 python scripts/profile_step.py 8 parity atrium 2 | grep Msam
timeout 300 python scripts/profile_step.py 8 parity cornell 2 | grep Msam
