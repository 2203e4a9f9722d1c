for m in 0 1 2; do
  NB200_CONV_PERSIST=$m timeout 300 python tools/conv_check.py > gpurun_out/conv_check_$m.log 2>&1; echo "mode $m rc=$?"; tail -12 gpurun_out/conv_check_$m.log
done
for m in 0 1 2; do
  NB200_CONV_PERSIST=$m timeout 300 python tools/conv_check.py c4 > gpurun_out/conv_c4_$m.log 2>&1; echo "c4 mode $m rc=$?"; tail -4 gpurun_out/conv_c4_$m.log
done
