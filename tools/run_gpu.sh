set -x
timeout 1500 python -m pytest tests/test_gpu_fullsize.py -x -q -s -k "oracle or band" > gpurun_out/pytest_full.log 2>&1; tail -15 gpurun_out/pytest_full.log
