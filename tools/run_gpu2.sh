timeout 900 python -m pytest tests/test_gpu_multidevice.py -x -q > gpurun_out/pytest_md.log 2>&1; tail -15 gpurun_out/pytest_md.log
cd netsdb_b200/host/bin
for nd in 1 2; do echo "== FFTestWithoutDeduplication_netsdb Amazon-14K NB200_DEVICES=$nd"; NB200_DEVICES=$nd NB200_DROPOUT_AS_ZERO=1 NB200_PRECISION=f16 timeout 900 ./FFTestWithoutDeduplication_netsdb 1000 1000 10000 597540 1000 14588 ff Y 2>&1 | tail -4; done > /root/repo/gpurun_out/ref_driver_multi.log 2>&1
cd /root/repo; cat gpurun_out/ref_driver_multi.log
