timeout 2400 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; tail -6 gpurun_out/pytest_gpu_full.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; python3 -c "
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['breakdown_ms'])"
