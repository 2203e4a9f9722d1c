cd netsdb_b200/host/bin
for nd in 1 8; do NB200_DEVICES=$nd NB200_PRECISION=f16 timeout 600 ./FFTest 1000 1000 4000 8200 1000 64 /tmp/dump$nd.txt > /tmp/fftest$nd.log 2>&1; echo "FFTest NB200_DEVICES=$nd rc=$?"; tail -2 /tmp/fftest$nd.log; done
python3 - <<'PY'
import numpy as np
a=np.loadtxt('/tmp/dump1.txt'); b=np.loadtxt('/tmp/dump8.txt')
print('rows', a.shape, 'labels equal between 1 and 8 devices:', bool(np.array_equal(a[:,0],b[:,0]) and np.array_equal(a[:,1],b[:,1])), 'max prob diff', float(np.abs(a[:,2]-b[:,2]).max()))
PY
echo "== FFTestWithoutDeduplication_netsdb (unmodified reference driver) Amazon-14K NB200_DEVICES=8"
NB200_DEVICES=8 NB200_DROPOUT_AS_ZERO=1 NB200_PRECISION=f16 timeout 1200 ./FFTestWithoutDeduplication_netsdb 1000 1000 10000 597540 1000 14588 ff Y 2>&1 | tail -4
