set -x
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -c 2500 gpurun_out/bench_n2.json; tail -5 gpurun_out/bench_n2.err
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "k_sharded" > gpurun_out/pytest_ksh.log 2>&1; tail -5 gpurun_out/pytest_ksh.log
