timeout 900 python -m pytest tests -m gpu -q -x -s -k "light_count or too_many" 2>&1 | grep -v "^$" | tail -25 | cut -c1-300
