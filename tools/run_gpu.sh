timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -k "udf or goldens or network" > gpurun_out/pytest_udf.log 2>&1; tail -15 gpurun_out/pytest_udf.log
timeout 1500 python -m pytest tests/test_gpu_fullsize.py -x -q -k "word2vec or semantic" > gpurun_out/pytest_udf2.log 2>&1; tail -15 gpurun_out/pytest_udf2.log
