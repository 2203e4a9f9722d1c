for cfg in 5 6 7; do for co in 72 90; do echo "cfg $cfg carve $co: $(NB200_F64_CFG=$cfg NB200_F64_CARVEOUT=$co timeout 300 python tools/f64_skinny_bench.py 2>&1 | tail -1)"; done; done
