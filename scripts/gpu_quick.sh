for c in 0.15 0.3 0.45 0.8 1.2; do echo "TRI_COST=$c"; PTC_TRI_COST=$c timeout 300 python scripts/profile_step.py 8 fast atrium 2 | cut -c29-140; done
