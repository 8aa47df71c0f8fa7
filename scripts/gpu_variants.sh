set -x
for v in "" tie; do
  if [ -n "$v" ]; then export PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_$v; else unset PTC_LIB_DIR; fi
  echo "VARIANT=$v"
  for i in 1 2; do python scripts/profile_step.py 8 fast atrium 3 2>&1 | tail -1; done
  python scripts/profile_step.py 1 fast terrain 3 2>&1 | tail -1
  python scripts/profile_step.py 8 fast cornell 3 2>&1 | tail -1
  python scripts/profile_step.py 8 parity atrium 1 2>&1 | tail -1
done
export PTC_LIB_DIR=$PWD/opencl_pathtracer_b200/_lib_tie
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
