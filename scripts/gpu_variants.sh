set -x
for v in "" ni; do
  if [ -n "$v" ]; then export PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_$v; else unset PTC_LIB_DIR; fi
  echo "VARIANT=$v"
  for i in 1 2; do python scripts/profile_step.py 8 fast atrium 3 2>&1 | tail -1; done
  python scripts/profile_step.py 8 fast cornell 3 2>&1 | tail -1
done
