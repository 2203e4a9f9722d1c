"""Times config C4 (3x3 conv, 64->64 channels, 112x112, batch 64, + ReLU) through nb200_conv2d_fused.
The call takes HOST images (the reference's Image set) so the wall time includes H2D of 411 MB and the NCHW->NHWC
conversion; run under `ncu --metrics gpu__time_duration.sum` for the per-kernel device times."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from netsdb_b200 import capi

capi.init(0)
rng = np.random.default_rng(0)
n, c, h, w, kout, k = 64, 64, 112, 112, 64, 3
imgs = rng.uniform(-0.5, 0.5, size=(n, c, h, w))
kern = rng.uniform(-0.5, 0.5, size=(kout, c, k, k))
bias = rng.uniform(-0.5, 0.5, size=kout)
for prec in ("bf16", "bf16x3"):
    for it in range(4):
        t0 = time.perf_counter()
        R = capi.conv2d_fused(imgs, kern, bias, 0, True, capi.PRECISIONS[prec], 1000, 1000)
        capi.synchronize()
        dt = time.perf_counter() - t0
    rows = R.info()[0]
    print(f"C4 {prec}: {dt * 1e3:.2f} ms per call (host images in, result resident), {rows} output rows, "
          f"{2 * rows * kout * c * k * k / dt / 1e12:.2f} useful TFLOP/s incl. H2D", flush=True)
capi.shutdown()
