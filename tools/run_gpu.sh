for k in 0 1 2; do NB200_GEMM_KERNEL_SHORTK=$k timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python3 -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); b=d['breakdown_ms']; print('shortk=$k', d['ms_per_step'], b['ms_gemm2'], b['ms_softmax'], b['ms_refine'], d['strong_check'])"; done
