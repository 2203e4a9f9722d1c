"""Development check of the streamed layer-1 kernel (nb200_l1stream.cu) through the C ABI.

python tools/l1s_check.py parity     small/ragged shapes: nb200_ff_layer1_partial(F16) with the cluster kernel against
                                      numpy fp64 and against the separate conversion + GEMM kernels
python tools/l1s_check.py amazon     the Amazon-14K shape: layer-1 partial of both paths compared, then the stage timings
                                      of nb200_ff_inference_unit for fused_layer1 = 2 / 1
Environment (read once per process): NB200_L1S_MT=1|2, NB200_L1S_CLUSTERS=n, NB200_L1S_DEBUG=bits.
"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from netsdb_b200 import capi as nb  # noqa: E402


def partial(w1, x, rows, H):
    out = torch.zeros((rows, H + 1), dtype=torch.float32, device="cuda")
    nb.check(nb.lib().nb200_ff_layer1_partial(w1.ptr, x.ptr, nb.PREC_F16, C.c_void_p(out.data_ptr()), H + 1))
    nb.synchronize()
    torch.cuda.synchronize()
    return out.cpu().numpy()


def parity():
    bad = 0
    for rows, F, H in ((300, 8228, 1000), (1250, 20000, 1000), (700, 12000, 777), (2600, 9000, 1024), (5, 4100, 600)):
        w1 = nb.DeviceSet.create("t", "w1", H, F, 1000, 1000); w1.fill_synth(2)
        x = nb.DeviceSet.create("t", "x", rows, F, 1000, 1000); x.fill_synth(1)
        X, W = x.download_dense(), w1.download_dense()
        want = X @ W.T
        ssq = (X.astype(np.float32).astype(np.float64) ** 2).sum(axis=1)
        res = {}
        for mode in (2, 0):
            nb.set_option("fused_layer1", mode)
            x.invalidate_planes()
            n0 = nb.launch_count()
            got = partial(w1, x, rows, H)
            res[mode] = got
            err = np.abs(got[:, :H] - want).max() / np.abs(want).max()
            ess = np.abs(got[:, H] - ssq).max() / ssq.max()
            ok = err < 2.5e-3 and ess < 1e-4
            bad += not ok
            print(f"rows={rows} F={F} H={H} mode={mode} launches={nb.launch_count() - n0} rel_err={err:.2e} sumsq_err={ess:.2e} {'OK' if ok else 'FAIL'}",
                  flush=True)
        d = np.abs(res[2][:, :H] - res[0][:, :H]).max() / np.abs(want).max()
        print(f"   stream vs separate kernels: {d:.2e}", flush=True)
        bad += not d < 1e-5
        w1.destroy(); x.destroy()
    print("PARITY", "FAIL" if bad else "OK", flush=True)
    return bad


def amazon(batch=10000):
    feat, hid, lab, blk = 597540, 1000, 14588, 1000
    DS = nb.DeviceSet
    w1 = DS.create("ff", "w1", hid, feat, blk, blk); w1.fill_synth(2)
    b1 = DS.create("ff", "b1", hid, 1, blk, 1); b1.fill_synth(3)
    wo = DS.create("ff", "wo", lab, hid, blk, blk); wo.fill_synth(4)
    bo = DS.create("ff", "bo", lab, 1, blk, 1); bo.fill_synth(5)
    x = DS.create("ff", "inputs", batch, feat, blk, blk); x.fill_synth(1)
    labels = {}
    for mode in ((2,) if os.environ.get("L1S_ONLY") else (2, 1)):
        nb.set_option("fused_layer1", mode)
        for rep in range(3 if os.environ.get("L1S_ONLY") else 4):
            x.invalidate_planes()
            _, lab_, st = nb.ff_inference_unit(w1, wo, x, b1, bo, nb.PREC_F16)
            d = st if isinstance(st, dict) else st.as_dict()
            print(f"mode={mode} rep={rep} " + " ".join(f"{k}={v:.3f}" if isinstance(v, float) else f"{k}={v}" for k, v in d.items()), flush=True)
        labels[mode] = lab_
    if 1 in labels:
        print("labels equal between modes:", bool(np.array_equal(labels[2], labels[1])), flush=True)
    if batch <= 2000:
        res = {}
        for mode in (2, 0):
            nb.set_option("fused_layer1", mode)
            x.invalidate_planes()
            res[mode] = partial(w1, x, batch, hid)
        s = np.abs(res[0][:, :hid]).max()
        print(f"partial stream vs separate: {np.abs(res[2][:, :hid] - res[0][:, :hid]).max() / s:.2e}", flush=True)


if __name__ == "__main__":
    nb.init(0)
    what = sys.argv[1] if len(sys.argv) > 1 else "parity"
    rc = 0
    if what == "parity":
        rc = parity()
    else:
        amazon(int(sys.argv[2]) if len(sys.argv) > 2 else 10000)
    nb.shutdown()
    sys.exit(1 if rc else 0)
