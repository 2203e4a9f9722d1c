"""torchrun script: K-sharded FF inference on WORLD_SIZE GPUs against the CPU oracle (rank 0 checks).
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/tools/ksharded_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from netsdb_b200 import capi  # noqa: E402
from netsdb_b200.ksharded import col_shard, ff_inference_k_sharded  # noqa: E402
from oracle import ff_oracle as fo  # noqa: E402  (the checker; this is a test tool)

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
capi.init(local)
capi.set_stream(torch.cuda.current_stream().cuda_stream)
batch, feat, hid, lab, blk = 1003, 9000, 500, 300, 500
x, w1 = fo.synth_dense(batch, feat, 1), fo.synth_dense(hid, feat, 2)
b1, wo, bo = fo.synth_dense(hid, 1, 3), fo.synth_dense(lab, hid, 4), fo.synth_dense(lab, 1, 5)
c0, nc = col_shard(feat, blk, world, rank)
DS = capi.DeviceSet
X = DS.create("ff", "inputs", batch, nc, blk, blk); X.upload_dense(np.ascontiguousarray(x[:, c0:c0 + nc]))
W1 = DS.create("ff", "w1", hid, nc, blk, blk); W1.upload_dense(np.ascontiguousarray(w1[:, c0:c0 + nc]))
B1 = DS.create("ff", "b1", hid, 1, blk, 1); B1.upload_dense(b1)
WO = DS.create("ff", "wo", lab, hid, blk, blk); WO.upload_dense(wo)
BO = DS.create("ff", "bo", lab, 1, blk, 1); BO.upload_dense(bo)
ok = True
for prec, tol in (("f16", 2.5e-3), ("bf16x3", 1e-4), ("bf16", 2e-2)):
    out, labels, undecided, row0 = ff_inference_k_sharded(dist, torch, rank, world, W1, X, B1, WO, BO, capi.PRECISIONS[prec])
    z = DS.lookup("ff", "__logits").download_dense()
    gathered = [None] * world
    dist.all_gather_object(gathered, (row0, labels, z, undecided))
    if rank == 0:
        L = lambda d, bc: fo.load_matrix(d, blk, bc)
        _, ref_logits, ref_labels = fo.ff_inference_unit(L(w1, blk), L(wo, blk), L(x, blk), L(b1, 1), L(bo, 1), threads=4)
        zr = ref_logits.dense()
        all_labels = np.concatenate([g[1] for g in sorted(gathered, key=lambda g: g[0])])
        all_z = np.concatenate([g[2] for g in sorted(gathered, key=lambda g: g[0])])
        err = np.abs(all_z - zr).max() / np.abs(zr).max()
        nund = sum(g[3] for g in gathered)
        nbad = int((all_labels != ref_labels).sum())
        good = err <= tol and nbad <= nund
        ok &= good
        print(f"{'OK ' if good else 'BAD'} k-sharded world={world} {prec}: logit rel err {err:.2e} (tol {tol}), "
              f"labels differing {nbad} of {batch} (undecided rows reported: {nund})", flush=True)
capi.shutdown()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
