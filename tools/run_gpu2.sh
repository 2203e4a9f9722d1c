timeout 900 python -m pytest tests/test_gpu_multidevice.py tests/test_gpu_housekeeping.py tests/test_gpu_l1stream.py -x -q > gpurun_out/pytest_md.log 2>&1; tail -15 gpurun_out/pytest_md.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_lstm.py tests/test_la.py tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_md2.log 2>&1; tail -5 gpurun_out/pytest_md2.log
