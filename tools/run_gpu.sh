timeout 2400 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; tail -6 gpurun_out/pytest_gpu_full.log
timeout 900 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -12 gpurun_out/smoke.log
timeout 1500 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 600 gpurun_out/bench_n1.json; tail -3 gpurun_out/bench_n1.err
