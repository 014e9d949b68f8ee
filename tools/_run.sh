timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for w in config2 headline config3; do
python bench.py --workload $w --steps 10 --warmup 3 --skip-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w',round(d['value']),round(d['ms_per_step'],3),d['stage_ms'])
"
done
python tools/_tl.py 2>&1 | tail -8
