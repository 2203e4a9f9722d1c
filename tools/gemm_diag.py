"""GPU diagnostic for the tcgen05 block GEMM: small integer-valued operands (exact in bf16, exact
fp32 sums) against numpy.  Prints a pass/fail line per shape; exits non-zero on any mismatch."""
import sys
import numpy as np

sys.path.insert(0, ".")
from netsdb_b200 import capi


def run(M, N, K, op, prec, bm=100, bn=100, bk=100):
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    a = rng.integers(-3, 4, size=(M, K)).astype(np.float64)
    if op == capi.OP_NT:
        b = rng.integers(-3, 4, size=(N, K)).astype(np.float64)
        ref = a @ b.T
        B = capi.DeviceSet.create("diag", "b", N, K, bn, bk)
    else:
        b = rng.integers(-3, 4, size=(K, N)).astype(np.float64)
        ref = a @ b
        B = capi.DeviceSet.create("diag", "b", K, N, bk, bn)
    A = capi.DeviceSet.create("diag", "a", M, K, bm, bk)
    A.upload_dense(a)
    B.upload_dense(b)
    C = capi.gemm_blocks(op, A, B, capi.EPI_NONE, None, prec, "diag", "c")
    got = C.download_dense()
    err = np.abs(got - ref).max()
    bad = np.argwhere(np.abs(got - ref) > 1e-6)
    tag = "OK " if err == 0 else "BAD"
    print(f"{tag} M={M} N={N} K={K} op={op} prec={prec} maxerr={err:.3g} nbad={len(bad)}"
          + (f" first_bad={bad[:4].tolist()} got={got[tuple(bad[0])]:.4g} ref={ref[tuple(bad[0])]:.4g}" if len(bad) else ""),
          flush=True)
    return err == 0


if __name__ == "__main__":
    capi.init(0)
    ok = True
    shapes = [(128, 256, 64), (128, 256, 256), (100, 200, 130), (300, 700, 1000), (128, 64, 64), (257, 129, 4160),
              (1000, 1100, 20000)]
    for prec in (capi.PREC_FP64, capi.PREC_BF16, capi.PREC_BF16X3):
        for (M, N, K) in shapes:
            for op in (capi.OP_NT, capi.OP_NN):
                try:
                    ok &= run(M, N, K, op, prec)
                except Exception as e:  # keep going: one report covers every failure
                    print(f"EXC M={M} N={N} K={K} op={op} prec={prec}: {e}", flush=True)
                    ok = False
    capi.shutdown()
    sys.exit(0 if ok else 1)
