timeout 300 python tools/conv_check.py 2>&1 | tail -10
for d in 0 2; do
  NB200_CONV_DEBUG=$d timeout 300 python tools/conv_check.py c4 > gpurun_out/conv_c4_d$d.log 2>&1; echo "c4 debug $d rc=$?"; grep -a "kernel_ms\|total" gpurun_out/conv_c4_d$d.log | tail -4
done
