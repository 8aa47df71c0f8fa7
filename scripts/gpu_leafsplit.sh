set -x
for v in 0 1; do for i in 1 2; do PTC_LEAF_SPLIT=$v python scripts/profile_step.py 8 fast atrium 3 2>&1 | tail -1; done; done
for v in 0 1; do PTC_LEAF_SPLIT=$v python scripts/profile_step.py 1 fast terrain 3 2>&1 | tail -1; done
for v in 0 1; do PTC_LEAF_SPLIT=$v python scripts/profile_step.py 8 fast cornell 3 2>&1 | tail -1; done
