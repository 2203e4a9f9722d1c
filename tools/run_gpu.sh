timeout 900 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python3 -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print(d['value'], d['ms_per_step'], r['frac'], r['frac_sustained'], r['peak_sustained'], r['traffic'], r['hbm']['frac'])"
timeout 300 python bench.py --config c1 --steps 20 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | python3 -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c1', d['value'], d['ms_per_step'], d['config']['workload'])"
