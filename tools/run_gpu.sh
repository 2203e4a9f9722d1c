timeout 2400 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; tail -4 gpurun_out/pytest_gpu_full.log
timeout 900 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
