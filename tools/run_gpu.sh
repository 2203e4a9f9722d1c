export NB200_L1S_VERBOSE=1 L1S_ONLY=1
{
echo "=== parity"; timeout 300 python tools/l1s_check.py parity 2>&1 | grep -E "PARITY|rror|fault" | tail -3
for dual in 1 0 1 0; do
  echo "=== amazon DUAL=$dual"
  NB200_L1S_DUAL=$dual timeout 300 python tools/l1s_check.py amazon 2>&1 | grep -E "l1 stream\]|rep=1|rep=2|rror|fault" | sed -E 's/ms_convert.*ms_gemm1=/g1=/; s/ms_gemm2.*ms_gemm1_kernel/k1/; s/refined.*//' | tail -5
done
echo "=== 5000 rows"
for dual in 1 0; do NB200_L1S_DUAL=$dual timeout 300 python tools/l1s_check.py amazon 5000 2>&1 | grep -E "l1 stream\]|rep=2|rror|fault" | sed -E 's/ms_convert.*ms_gemm1=/g1=/; s/ms_gemm2.*ms_gemm1_kernel/k1/; s/refined.*//' | tail -3; done
} > gpurun_out/l1s_dual.log 2>&1
timeout 600 python -m pytest tests/test_gpu_l1stream.py tests/test_gpu_parity.py -x -q -k "stream or fused_layer1 or config1" > gpurun_out/pytest_dual.log 2>&1; tail -3 gpurun_out/pytest_dual.log
