"""Times the last cascade level's layer-1 product in isolation: a few fp64 rows against all of W1 (1000 x 597,540),
nb200_gemm_blocks(NT, FP64).  NB200_F64_CFG / NB200_F64_SKINNY choose the kernel variant."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from netsdb_b200 import capi

rows, K, N = (int(sys.argv[1]) if len(sys.argv) > 1 else 15), 597540, 1000
capi.init(0)
torch.cuda.set_device(0)
stream = torch.cuda.Stream()
capi.set_stream(stream.cuda_stream)
A = capi.DeviceSet.create("t", "a", rows, K, 1000, 1000); A.fill_synth(1)
W = capi.DeviceSet.create("t", "w", N, K, 1000, 1000); W.fill_synth(2)
ts = []
for it in range(8):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        C = capi.gemm_blocks(capi.OP_NT, A, W, capi.EPI_NONE, None, capi.PREC_FP64, "t", "c")
        e1.record(stream)
    e1.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = float(np.median(ts[3:]))
print(f"fp64 skinny {rows} x {N} x {K}: {ms:.3f} ms  ({N * K * 8 / ms / 1e6:.0f} GB/s of W1, {2 * 16 * 1024 * K / ms / 1e9:.1f} TFLOP/s executed)  checksum {C.download_dense().sum():.6e}")
capi.shutdown()
