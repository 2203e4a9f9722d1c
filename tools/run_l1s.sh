export NB200_L1S_VERBOSE=1 L1S_ONLY=1
{
for mt in 2 1; do
  echo "=== parity MT=$mt"
  NB200_L1S_MT=$mt timeout 300 python tools/l1s_check.py parity 2>&1 | grep -E "mode=2|PARITY|rror|fault" | tail -4
done
for cfg in "2 0 64" "2 0 45" "2 2 64" "2 4 64" "2 6 64" "1 0 64" "1 0 45"; do
  set -- $cfg
  echo "=== amazon MT=$1 DEBUG=$2 RING=$3"
  NB200_L1S_MT=$1 NB200_L1S_DEBUG=$2 NB200_L1S_RING=$3 timeout 300 python tools/l1s_check.py amazon 2>&1 | grep -E "rep=2|rror|fault" | sed -E 's/ms_convert.*ms_gemm1_kernel/k1/; s/refined.*//'
done
} > gpurun_out/l1s_v7.log 2>&1
