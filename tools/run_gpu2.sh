timeout 1200 python -m pytest tests/test_gpu_multidevice.py tests/test_gpu_parity.py -x -q -k "multidevice or two_gpus or devices or k_sharded" > gpurun_out/pytest_2gpu.log 2>&1; tail -6 gpurun_out/pytest_2gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -c 400 gpurun_out/bench_n2.err; python3 -c "
import json
d=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','scaling','n_gpus')}); print('e2e', d['e2e']['value'] if d['e2e'] else None); print(d['weak_replicas']); print(d['strong_check']); print(d['ksharded']); print(d['breakdown_ms'])"
