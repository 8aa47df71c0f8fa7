timeout 900 python -m pytest tests -m gpu -q -x -s -k "device_built or gpu_bvh_flag" 2>&1 | grep -v "^$" | tail -30 | cut -c1-300
