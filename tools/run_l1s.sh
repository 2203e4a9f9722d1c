export NB200_L1S_VERBOSE=1 L1S_ONLY=1
{
for relay in 0 1; do for mt in 2 1; do
  echo "=== parity MT=$mt RELAY=$relay"
  NB200_L1S_RELAY=$relay NB200_L1S_MT=$mt timeout 300 python tools/l1s_check.py parity 2>&1 | grep -E "mode=2|PARITY|rror|fault"
done; done
for cfg in "2 0 0 0" "2 1 0 0" "1 0 0 0" "2 0 1 0" "2 0 2 0" "2 0 4 0" "2 0 0 8" "2 0 0 16" "1 0 0 8" "2 1 0 8"; do
  set -- $cfg
  echo "=== amazon MT=$1 RELAY=$2 DEBUG=$3 PF=$4"
  NB200_L1S_MT=$1 NB200_L1S_RELAY=$2 NB200_L1S_DEBUG=$3 NB200_L1S_PF=$4 timeout 300 python tools/l1s_check.py amazon 2>&1 | grep -E "rep=2|rror|fault" | sed -E 's/ms_convert.*ms_gemm1_kernel/k1/; s/refined.*//'
done
} > gpurun_out/l1s_v2.log 2>&1
