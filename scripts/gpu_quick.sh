timeout 900 python -m pytest tests -m gpu -q -s -k "bounce_like or wide_bvh_hits or fast_mode" 2>&1 | grep "vs PARITY\|farther\|mismatch\|passed\|failed\|wide vs" | grep -v "print\|assert" | cut -c1-300
timeout 300 python scripts/profile_step.py 8 fast atrium 2
