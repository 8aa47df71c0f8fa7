timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for nm in 2 3 4; do for tm in 2 4 6; do echo "NODE_MIN=$nm TRI_MIN=$tm"; PTC_NODE_MIN=$nm PTC_TRI_MIN=$tm timeout 300 python scripts/profile_step.py 4 fast atrium 2; done; done
