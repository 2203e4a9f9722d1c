"""Random-shape sweep of nb200_gemm_blocks (all join flavours, epilogues and precisions) against numpy fp64.
Prints one line per failure and a summary; exit status 1 if anything failed."""
import sys

import numpy as np

sys.path.insert(0, ".")
from netsdb_b200 import capi

TOL = {"fp64": 1e-12, "bf16x3": 1e-4, "f16": 2.5e-3, "bf16": 2e-2}
n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 150
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
capi.init(0)
fails = 0
for case in range(n_cases):
    big = rng.random() < 0.25
    I = int(rng.integers(1, 1400 if big else 300))
    J = int(rng.integers(1, 1400 if big else 300))
    K = int(rng.integers(1, 9000 if big else 700))
    Kb = K if rng.random() < 0.8 else int(rng.integers(1, K + 1))  # differing K extents: contraction over min
    bi, bj, bk = (int(rng.integers(1, 200)) for _ in range(3))
    op = int(rng.integers(0, 3))
    prec = ["fp64", "bf16x3", "f16", "bf16"][int(rng.integers(0, 4))]
    epi = int(rng.integers(0, 3)) if op != capi.OP_TN else 0
    a = rng.uniform(-1, 1, size=(K, I) if op == capi.OP_TN else (I, K))
    b = rng.uniform(-1, 1, size=(J, Kb) if op == capi.OP_NT else (Kb, J))
    bias = rng.uniform(-1, 1, size=(I, 1))
    km = min(K, Kb)
    am = a[:km].T if op == capi.OP_TN else a[:, :km]
    bm = b[:, :km].T if op == capi.OP_NT else b[:km]
    want = am @ bm
    if epi == capi.EPI_BIAS_RELU:
        want = np.maximum(want + bias, 0)
    elif epi == capi.EPI_BIAS_TRANSPOSE:
        want = (want + bias).T
    try:
        A = capi.DeviceSet.create("fz", "a", a.shape[0], a.shape[1], bk if op == capi.OP_TN else bi, bi if op == capi.OP_TN else bk)
        B = capi.DeviceSet.create("fz", "b", b.shape[0], b.shape[1], bj if op == capi.OP_NT else bk, bk if op == capi.OP_NT else bj)
        A.upload_dense(a); B.upload_dense(b)
        Bs = None
        if epi:
            Bs = capi.DeviceSet.create("fz", "bias", I, 1, bi, 1); Bs.upload_dense(bias)
        C = capi.gemm_blocks(op, A, B, epi, Bs, capi.PRECISIONS[prec], "fz", "c")
        got = C.download_dense()
        # error scale of a tensor-core product: |a| |b| summed, not the (possibly cancelling) result
        scale = max(np.abs(am).__matmul__(np.abs(bm)).max(), 1e-300) if prec != "fp64" else max(np.abs(want).max(), 1e-300)
        err = np.abs(got - want).max() / scale
        ok = got.shape == want.shape and err <= TOL[prec]
    except Exception as e:  # noqa: BLE001
        ok, err = False, repr(e)
    if not ok:
        fails += 1
        print(f"FAIL case {case}: op {op} epi {epi} {prec} I {I} J {J} K {K}/{Kb} blocks {bi},{bj},{bk}: {err}", flush=True)
print(f"{n_cases - fails}/{n_cases} OK")
capi.shutdown()
sys.exit(1 if fails else 0)
