timeout 1200 python -m pytest tests/test_gpu_multidevice.py -x -q > gpurun_out/pytest_2gpu.log 2>&1; tail -6 gpurun_out/pytest_2gpu.log
