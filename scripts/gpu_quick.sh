timeout 900 python -m pytest tests -m gpu -q -x -s -k "device_built" 2>&1 | grep -v "^$" | tail -12 | cut -c1-200
PTC_GPU_BVH=1 timeout 300 python scripts/profile_step.py 8 fast atrium 2
PTC_GPU_BVH=1 PTC_GPU_BVH_BUILDER=lbvh timeout 300 python scripts/profile_step.py 8 fast atrium 2
PTC_GPU_BVH=1 timeout 600 python scripts/profile_step.py 2 fast terrain 2
