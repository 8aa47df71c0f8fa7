PTC_GPU_BVH=1 timeout 300 python scripts/profile_step.py 8 fast atrium 2
PTC_GPU_BVH=1 timeout 600 python scripts/profile_step.py 2 fast terrain 2
