timeout 900 python -m pytest tests -m gpu -q -x -s -k "bilinear" 2>&1 | grep -i "bilinear:\|passed\|failed\|^E " | head
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
