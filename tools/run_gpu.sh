# launch list of the default bench command (short: no e2e, no CPU leg) -- per-launch times are cold-cache and serialised
NB200_WAIT_LOG2=38 timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; tail -2 gpurun_out/ncu_launch.log | cut -c1-300
# full capture of the layer-1 kernel (third launch: warm) and of the conv kernel
NB200_WAIT_LOG2=38 L1S_ONLY=1 timeout 1500 ncu --set full --import-source on --clock-control none -k regex:k_l1_stream -s 2 -c 1 -o gpurun_out/r02_l1stream python tools/l1s_check.py amazon > gpurun_out/ncu_l1s.log 2>&1; tail -3 gpurun_out/ncu_l1s.log
NB200_WAIT_LOG2=38 timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_conv_persist -s 3 -c 1 -o gpurun_out/r02_conv_persist python tools/conv_check.py c4 > gpurun_out/ncu_conv.log 2>&1; tail -3 gpurun_out/ncu_conv.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 700 gpurun_out/bench_reference.json
