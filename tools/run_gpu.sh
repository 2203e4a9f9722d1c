timeout 900 python -m pytest tests/test_gpu_l1stream.py tests/test_gpu_housekeeping.py -x -q > gpurun_out/pytest_l1s.log 2>&1; tail -5 gpurun_out/pytest_l1s.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -8 gpurun_out/smoke.log
timeout 900 python bench.py --config c5 --steps 5 --warmup 3 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; tail -c 600 gpurun_out/bench_c5.json; tail -3 gpurun_out/bench_c5.err
