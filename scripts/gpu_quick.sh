for c in 50 66 75 100; do echo "CARVEOUT=$c"; PTC_CARVEOUT=$c timeout 300 python scripts/profile_step.py 8 fast atrium 2 | grep Msam | cut -c29-140; done
