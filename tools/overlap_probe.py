"""Does the next batch's upload really run under the current batch's computation?  Queues the 47.8 GB upload of set B,
then times (host wall clock) the inference on set A and the download of its output."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from netsdb_b200 import capi

batch, feat, hid, lab, blk = 10000, 597540, 1000, 14588, 1000
capi.init(0)
DS = capi.DeviceSet
w1 = DS.create("ff", "w1", hid, feat, blk, blk); w1.fill_synth(2)
b1 = DS.create("ff", "b1", hid, 1, blk, 1); b1.fill_synth(3)
wo = DS.create("ff", "wo", lab, hid, blk, blk); wo.fill_synth(4)
bo = DS.create("ff", "bo", lab, 1, blk, 1); bo.fill_synth(5)
A = DS.create("ff", "inputs", batch, feat, blk, blk); A.fill_synth(1)
B = DS.create("ff", "inputs_b", batch, feat, blk, blk); B.fill_synth(7)
capi.synchronize()
nblk = 64  # reuse one pinned 8 MB block for every upload: only the transfer time matters here
host = torch.empty(blk * blk, dtype=torch.float64).pin_memory().numpy().reshape(blk, blk)
host[:] = 0.25
out_host = torch.empty(batch * lab, dtype=torch.float64).pin_memory().numpy().reshape(batch, lab)
prec = capi.PRECISIONS["f16"]
for _ in range(2):
    capi.ff_inference_unit(w1, wo, A, b1, bo, prec)
n_rb, n_cb = batch // blk, (feat + blk - 1) // blk


def upload(dst):
    t = time.perf_counter()
    for i in range(n_rb):
        for j in range(n_cb):
            cols = min(blk, feat - j * blk)
            capi.check(capi.lib().nb200_upload_block(dst.ptr, i, j, host.ctypes.data, blk, cols) if cols == blk else
                       capi.lib().nb200_upload_rect(dst.ptr, i * blk, j * blk, host.ctypes.data, blk, cols, blk))
    dst.upload_commit()
    return time.perf_counter() - t


A.upload_commit()
for rep in range(3):
    t0 = time.perf_counter()
    tu = upload(B)
    t1 = time.perf_counter()
    out, labels, st = capi.ff_inference_unit(w1, wo, A, b1, bo, prec)
    t2 = time.perf_counter()
    capi.check(capi.lib().nb200_download_rect(out.ptr, 0, 0, out_host.ctypes.data, batch, lab, lab))
    t3 = time.perf_counter()
    capi.synchronize()
    t4 = time.perf_counter()
    print(f"rep {rep}: enqueue upload {tu * 1e3:.1f} ms, inference call {1e3 * (t2 - t1):.1f} ms (device {st['ms_total']:.1f}), "
          f"download {1e3 * (t3 - t2):.1f} ms, drain {1e3 * (t4 - t3):.1f} ms, total {1e3 * (t4 - t0):.1f} ms", flush=True)
capi.shutdown()
