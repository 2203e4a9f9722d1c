for d in 0 1 2; do NB200_F64_TMA_DEBUG=$d timeout 300 python tools/f64_skinny_bench.py 2>&1 | tail -1; done
