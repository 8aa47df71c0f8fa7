timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
timeout 900 python -m pytest tests -m gpu -q -s -k "bounce_like or wide_bvh_hits or fast_mode" 2>&1 | grep "vs PARITY\|farther\|mismatch\|passed\|failed\|wide vs" | cut -c1-600
