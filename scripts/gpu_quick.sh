timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python scripts/profile_step.py 4 fast atrium 2
timeout 600 python bench.py --workload cornell --progressive 600 2>&1 | tail -3
timeout 600 python bench.py --workload atrium --progressive 200 2>&1 | tail -1
timeout 900 python bench.py --workload terrain --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['mrays_per_s'], d['stage_ms_per_step'], d['e2e'])"
