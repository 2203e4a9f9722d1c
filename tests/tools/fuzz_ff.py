"""Random-shape sweep of nb200_ff_inference_unit (all precisions; shapes on both sides of the fused-kernel and split-K
thresholds) against a dense numpy fp64 evaluation: labels must be identical, logits inside the precision's bound."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from netsdb_b200 import capi
from oracle import ff_oracle as fo  # checker only

TOL = {"fp64": 1e-12, "bf16x3": 1e-4, "f16": 2.5e-3, "bf16": 2e-2}
n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
capi.init(0)
fails = 0
for case in range(n_cases):
    big = rng.random() < 0.4
    batch = int(rng.integers(1, 3200 if big else 400))
    feat = int(rng.integers(1, 9500 if big else 900))
    hid = int(rng.integers(1, 1100 if big else 300))
    lab = int(rng.integers(1, 300))
    blk = int(rng.integers(20, 1000))
    prec = ["fp64", "bf16x3", "f16", "bf16"][int(rng.integers(0, 4))]
    d = dict(x=fo.synth_dense(batch, feat, 1 + case), w1=fo.synth_dense(hid, feat, 2), b1=fo.synth_dense(hid, 1, 3),
             wo=fo.synth_dense(lab, hid, 4), bo=fo.synth_dense(lab, 1, 5))
    z = (np.maximum(d["w1"] @ d["x"].T + d["b1"], 0).T @ d["wo"].T + d["bo"].T)
    want = z.argmax(axis=1).astype(np.int32)
    try:
        dims = dict(x=(batch, feat, blk, blk), w1=(hid, feat, blk, blk), b1=(hid, 1, blk, 1), wo=(lab, hid, blk, blk), bo=(lab, 1, blk, 1))
        dev = {}
        for k, (r, c, br, bc) in dims.items():
            dev[k] = capi.DeviceSet.create("fz", k, r, c, br, bc)
            dev[k].upload_dense(d[k])
        out, labels, st = capi.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], capi.PRECISIONS[prec], "fz", "output")
        zz = capi.DeviceSet.lookup("fz", "__logits").download_dense()
        err = np.abs(zz - z).max() / max(np.abs(z).max(), 1e-300)
        p = out.download_dense()
        ok = np.array_equal(labels, want) and err <= TOL[prec] and np.abs(p.sum(axis=1) - 1).max() < 1e-9
        msg = f"labels differ on {(labels != want).sum()} rows, logit err {err:.2e}, refined {st['refined_rows']}/{st['relevel_rows']}"
    except Exception as e:  # noqa: BLE001
        ok, msg = False, repr(e)
    if not ok:
        fails += 1
        print(f"FAIL case {case}: {prec} batch {batch} feat {feat} hid {hid} lab {lab} blk {blk}: {msg}", flush=True)
print(f"{n_cases - fails}/{n_cases} OK")
capi.shutdown()
sys.exit(1 if fails else 0)
