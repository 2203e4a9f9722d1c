export NB200_L1S_VERBOSE=1 L1S_ONLY=1
{
for mt in 2 1; do
  echo "=== parity MT=$mt"
  NB200_L1S_MT=$mt timeout 300 python tools/l1s_check.py parity 2>&1 | grep -E "mode=2|PARITY|rror|fault" | tail -6
done
for cfg in "2 0" "2 0" "1 0"; do
  set -- $cfg
  echo "=== amazon MT=$1 DEBUG=$2"
  NB200_L1S_MT=$1 NB200_L1S_DEBUG=$2 timeout 300 python tools/l1s_check.py amazon 2>&1 | grep -E "rep=2|rror|fault|prof" | sed -E 's/ms_convert.*ms_gemm1_kernel/k1/; s/refined.*//' | tail -13
done
export NB200_L1S_PROF=1
for cfg in "2 0"; do
  set -- $cfg
  echo "=== PROF amazon MT=$1 DEBUG=$2"
  NB200_L1S_MT=$1 NB200_L1S_DEBUG=$2 timeout 300 python tools/l1s_check.py amazon 2>&1 | grep -E "rep=2|rror|fault|prof" | sed -E 's/ms_convert.*ms_gemm1_kernel/k1/; s/refined.*//' | tail -13
done
} > gpurun_out/l1s_v11.log 2>&1
