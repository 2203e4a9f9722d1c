timeout 600 python -m pytest tests/test_gpu_housekeeping.py -x -q > gpurun_out/pytest_hk.log 2>&1; tail -15 gpurun_out/pytest_hk.log
