"""Development check of the implicit-GEMM conv2d kernels through the C ABI against torch's fp64 convolution.

python tools/conv_check.py            shapes with padding, several 128-wide segments, >64 channels, 64/128/256 filters
python tools/conv_check.py c4         the conv2d_memory_fusion bench shape (64 x 64ch x 112 x 112, 3x3, 64 filters): kernel time
Environment (read once per process): NB200_CONV_PERSIST=0 (tile-per-CTA kernel) | 1 (persistent kernel) .
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from netsdb_b200 import capi as nb  # noqa: E402


def run(n, ch, h, w, kout, k, pad, relu, reps=1):
    rng = np.random.default_rng(n * 1000 + ch + h + kout + k)
    img = rng.standard_normal((n, ch, h, w))
    ker = rng.standard_normal((kout, ch, k, k)) / np.sqrt(ch * k * k)
    bias = rng.standard_normal(kout)
    want = torch.nn.functional.conv2d(torch.from_numpy(img).cuda(), torch.from_numpy(ker).cuda(), torch.from_numpy(bias).cuda(), padding=pad)
    if relu:
        want = want.clamp_min(0)
    want = want.permute(0, 2, 3, 1).reshape(-1, kout).cpu().numpy()
    s = nb.conv2d_images_create(img, "cc", "img")
    ms = []
    for _ in range(reps):
        r = nb.conv2d_fused_from_set(s, ker, bias, pad, relu, nb.PREC_BF16, 1000, 1000, "cc", "res")
        ms.append(nb.last_kernel_ms())
        got = r.download_dense()
        r.destroy()
    s.destroy()
    err = np.abs(got - want).max() / np.abs(want).max()
    ok = err < 1.5e-2
    print(f"n={n} ch={ch} {h}x{w} kout={kout} k={k} pad={pad} relu={relu}: rel_err={err:.2e} kernel_ms={min(ms):.4f} {'OK' if ok else 'FAIL'}",
          flush=True)
    return not ok


if __name__ == "__main__":
    nb.init(0)
    print("NB200_CONV_PERSIST =", os.environ.get("NB200_CONV_PERSIST", "(default)"), flush=True)
    bad = 0
    if len(sys.argv) > 1 and sys.argv[1] == "c4":
        bad += run(64, 64, 112, 112, 64, 3, 0, True, reps=6)
    else:
        for a in ((2, 64, 20, 30, 64, 3, 0, True), (3, 3, 17, 23, 8, 3, 1, False), (2, 64, 12, 200, 64, 3, 1, True),
                  (1, 100, 9, 140, 128, 5, 2, True), (2, 130, 10, 40, 200, 1, 0, False), (5, 16, 33, 31, 64, 7, 3, True),
                  (300, 8, 6, 6, 12, 3, 1, True), (2, 64, 40, 300, 256, 2, 0, False)):
            bad += run(*a)
    nb.shutdown()
    print("CONV", "FAIL" if bad else "OK", flush=True)
    sys.exit(1 if bad else 0)
