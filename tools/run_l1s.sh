set -x
export NB200_L1S_VERBOSE=1
NB200_L1S_MT=1 timeout 300 python tools/l1s_check.py parity > gpurun_out/l1s_parity_mt1.log 2>&1; echo "rc=$?" >> gpurun_out/l1s_parity_mt1.log
NB200_L1S_MT=2 timeout 300 python tools/l1s_check.py parity > gpurun_out/l1s_parity_mt2.log 2>&1; echo "rc=$?" >> gpurun_out/l1s_parity_mt2.log
tail -25 gpurun_out/l1s_parity_mt1.log gpurun_out/l1s_parity_mt2.log
NB200_L1S_MT=2 timeout 600 python tools/l1s_check.py amazon > gpurun_out/l1s_amazon_mt2.log 2>&1; echo "rc=$?" >> gpurun_out/l1s_amazon_mt2.log
NB200_L1S_MT=1 timeout 600 python tools/l1s_check.py amazon > gpurun_out/l1s_amazon_mt1.log 2>&1; echo "rc=$?" >> gpurun_out/l1s_amazon_mt1.log
tail -15 gpurun_out/l1s_amazon_mt2.log gpurun_out/l1s_amazon_mt1.log
