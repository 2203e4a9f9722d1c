"""GPU test of nb200_transpose_mult_noalloc, the entry point for callers that may not allocate device
memory (threads confined to a pre-sized HBM arena, code inside a stream capture, serving threads that
must not hit cudaMalloc's device-wide sync).  It keeps its scratch -- bf16 copies of the operands and
the split-K partial tiles -- in a fixed 64 MiB array inside the library instead of growing it with
cudaMalloc, so this test drives it with device buffers the caller (here: torch) already owns and checks
the result against the regular entry point, nb200_gemm_blocks, on the same inputs.

Collected last on purpose (file name sorts after the other tests): a run that stops at the first
failure has then already exercised everything else.  It also passes when run alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL_BF16 = 2e-2  # the single-pass bf16 bound the parity tests use (relative to max|result|)


def _call(nb, torch, a, b):
    """Run the allocation-free product on torch-owned device buffers; returns C as numpy."""
    ta = torch.from_numpy(a).cuda()
    tb = torch.from_numpy(b).cuda()
    tc = torch.full((a.shape[0], b.shape[0]), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    st = nb.lib().nb200_transpose_mult_noalloc(C.c_void_p(ta.data_ptr()), a.shape[1], C.c_void_p(tb.data_ptr()),
                                              b.shape[1], a.shape[0], b.shape[0], a.shape[1],
                                              C.c_void_p(tc.data_ptr()), b.shape[0], None)  # NULL = default stream
    nb.check(st)
    torch.cuda.synchronize()
    return tc.cpu().numpy()


def _regular(nb, a, b, blk):
    A = nb.DeviceSet.create("na", "a", a.shape[0], a.shape[1], blk, blk); A.upload_dense(a)
    B = nb.DeviceSet.create("na", "b", b.shape[0], b.shape[1], blk, blk); B.upload_dense(b)
    return nb.gemm_blocks(nb.OP_NT, A, B, nb.EPI_NONE, None, nb.PREC_BF16, "na", "c").download_dense()


def test_no_alloc_product_matches_regular_entry_point_twice_in_a_row(nb):
    import torch

    rng = np.random.default_rng(5)
    # two calls back to back on the same scratch: the second is smaller and must not see the first's leftovers
    for (I, J, K, blk) in [(300, 520, 4100, 100), (70, 33, 250, 50)]:
        a = rng.uniform(-0.5, 0.5, size=(I, K))
        b = rng.uniform(-0.5, 0.5, size=(J, K))
        ref = _regular(nb, a, b, blk)
        got = _call(nb, torch, a, b)
        assert np.isfinite(got).all()   # (a kernel-internal fault would have filled C with NaN)
        exact = a @ b.T                 # independent of the library: fp64 on the host
        assert np.abs(got - exact).max() / np.abs(exact).max() <= TOL_BF16, (I, J, K)
        err = np.abs(got - ref).max() / np.abs(ref).max()
        assert err <= TOL_BF16, (I, J, K, err)
        # same arithmetic (one bf16 pass, fp32 accumulate): the two paths differ only in summation order
        assert err <= 1e-5, err
        # and a repeat of the very same call reproduces its own result exactly
        assert np.array_equal(_call(nb, torch, a, b), got)


def test_no_alloc_padded_pitches_on_a_side_stream(nb):
    """Operands and result that are windows of larger device buffers (lda, ldb, ldc wider than the logical
    matrices), queued on a stream of the caller's; J <= 128 takes the narrow-tile kernel."""
    import torch

    rng = np.random.default_rng(9)
    I, J, K = 200, 96, 700
    big_a = rng.uniform(-0.5, 0.5, size=(I, K + 13))
    big_b = rng.uniform(-0.5, 0.5, size=(J, K + 40))
    ta, tb = torch.from_numpy(big_a).cuda(), torch.from_numpy(big_b).cuda()
    tc = torch.full((I, J + 7), float("nan"), dtype=torch.float64, device="cuda")
    side = torch.cuda.Stream()
    torch.cuda.synchronize()
    for _ in range(2):
        nb.check(nb.lib().nb200_transpose_mult_noalloc(C.c_void_p(ta.data_ptr()), K + 13, C.c_void_p(tb.data_ptr()), K + 40,
                                                      I, J, K, C.c_void_p(tc.data_ptr()), J + 7, C.c_void_p(side.cuda_stream)))
        side.synchronize()
        got = tc.cpu().numpy()
        exact = big_a[:, :K] @ big_b[:, :K].T
        assert np.abs(got[:, :J] - exact).max() / np.abs(exact).max() <= TOL_BF16
        assert np.isnan(got[:, J:]).all()   # nothing written beyond the logical width


def test_no_alloc_rejects_sizes_beyond_the_fixed_scratch(nb):
    import torch

    t = torch.zeros(8, dtype=torch.float64, device="cuda")
    p = C.c_void_p(t.data_ptr())
    # 20,000 x 20,000 x 4,096 would need ~1.9 GB of scratch: refused up front, nothing is launched
    st = nb.lib().nb200_transpose_mult_noalloc(p, 4096, p, 4096, 20000, 20000, 4096, p, 20000, None)
    assert st == 1
    msg = nb.lib().nb200_last_error().decode()
    assert "scratch" in msg and str(64 << 20) in msg
