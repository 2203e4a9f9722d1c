for r in 3 8 9 15; do echo "rows $r: $(timeout 300 python tools/f64_skinny_bench.py $r 2>&1 | tail -1)"; done
timeout 2400 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; tail -4 gpurun_out/pytest_gpu_full.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python3 -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['breakdown_ms'])"
