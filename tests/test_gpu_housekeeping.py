"""GPU tests of the boundary's housekeeping: the pinned buffer pool (memory/source/SharedMem.cc:128 is what netsDB would hand
to nb200_pin_pool), rectangle uploads out of a wider host matrix, and a shutdown / init round trip."""
import ctypes as C
import mmap

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_pin_pool_over_a_shared_mapping(nb):
    """netsDB's buffer pool is ONE anonymous shared mapping carved into pages (SharedMem.cc:128); nb200_pin_pool page-locks
    such a range once, block uploads out of it and downloads into it then run as asynchronous DMA, and nb200_unpin_pool
    releases it."""
    pool_bytes = 64 << 20
    pool = mmap.mmap(-1, pool_bytes)                      # MAP_SHARED | MAP_ANONYMOUS, like the reference's pool
    base = C.addressof(C.c_char.from_buffer(pool))
    nb.check(nb.lib().nb200_pin_pool(C.c_void_p(base), C.c_size_t(pool_bytes)))
    import torch
    assert torch.cuda.is_available()
    at = C.c_int(0)
    rt = C.CDLL("libcudart.so.12")
    # the range is now page-locked: cudaHostGetFlags succeeds only on registered / pinned host memory
    flags = C.c_uint(0)
    assert rt.cudaHostGetFlags(C.byref(flags), C.c_void_p(base + 4096)) == 0
    rows, cols, blk = 700, 900, 300
    pages = np.frombuffer(pool, dtype=np.float64, count=rows * cols).reshape(rows, cols)
    rng = np.random.default_rng(3)
    pages[:] = rng.uniform(-1, 1, size=(rows, cols))
    s = nb.DeviceSet.create("pool", "m", rows, cols, blk, blk)
    for i in range(0, rows, blk):
        for j in range(0, cols, blk):
            r, c_ = min(blk, rows - i), min(blk, cols - j)
            nb.check(nb.lib().nb200_upload_rect(s.ptr, i, j, C.c_void_p(pages[i:, j:].ctypes.data), r, c_, cols))
    back = np.frombuffer(pool, dtype=np.float64, count=rows * cols, offset=rows * cols * 8).reshape(rows, cols)
    nb.check(nb.lib().nb200_download_rect(s.ptr, 0, 0, C.c_void_p(back.ctypes.data), rows, cols, cols))
    assert np.array_equal(back, pages)
    s.destroy()
    nb.synchronize()
    del pages, back
    nb.check(nb.lib().nb200_unpin_pool(C.c_void_p(base)))
    assert rt.cudaHostGetFlags(C.byref(flags), C.c_void_p(base + 4096)) != 0
    rt.cudaGetLastError()
    del at


def test_upload_rect_from_a_column_offset_view(nb):
    """A rectangle that is a window of a wider host matrix, large enough (>= 1 MiB) for the staged upload path: the copy
    must stop at the last element of the rectangle's last row (the buffer ends there)."""
    rows, cols, off = 600, 400, 37
    big = np.random.default_rng(4).uniform(-1, 1, size=(rows, cols + off))
    view = big[:, off:]                                     # ld = cols + off; the last row ends with the buffer
    s = nb.DeviceSet.create("rect", "m", rows, cols, 200, 200)
    s.upload_rect(0, 0, view)
    assert np.array_equal(s.download_dense(), view)
    # the same out of pinned memory, where reading past the end would fault instead of passing silently
    import torch
    t = torch.zeros(rows * (cols + off), dtype=torch.float64).pin_memory()
    pb = t.numpy().reshape(rows, cols + off)
    pb[:] = big
    s.upload_rect(0, 0, pb[:, off:])
    assert np.array_equal(s.download_dense(), view)
    s.destroy()


def test_shutdown_releases_the_fused_stage_and_init_starts_clean(oracle):
    """nb200_shutdown frees the fused stage's scratch: a later nb200_init + nb200_ff_inference_unit must not
    reuse pointers into a destroyed context."""
    from netsdb_b200 import capi as nb

    def run():
        nb.init(0)
        dense = dict(x=oracle.synth_dense(300, 4200, 1), w1=oracle.synth_dense(600, 4200, 2), b1=oracle.synth_dense(600, 1, 3),
                     wo=oracle.synth_dense(12, 600, 4), bo=oracle.synth_dense(12, 1, 5))
        dims = dict(x=(300, 4200, 500, 500), w1=(600, 4200, 500, 500), b1=(600, 1, 500, 1), wo=(12, 600, 500, 500), bo=(12, 1, 500, 1))
        dev = {}
        for k, (r, c_, br, bc) in dims.items():
            dev[k] = nb.DeviceSet.create("rt", k, r, c_, br, bc)
            dev[k].upload_dense(dense[k])
        labels = {p: nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PRECISIONS[p], "rt", "out")[1]
                  for p in ("f16", "bf16x3", "fp64")}
        assert np.array_equal(labels["f16"], labels["fp64"]) and np.array_equal(labels["bf16x3"], labels["fp64"])
        return labels["fp64"]

    a = run()
    nb.shutdown()
    b = run()
    assert np.array_equal(a, b)
    # (the session fixture `nb`, if another test created it, stays valid: init is idempotent for the same device)


def test_conv2d_on_a_resident_image_set_equals_the_host_image_call(nb, oracle):
    """nb200_conv2d_images_create + nb200_conv2d_fused_from_set (the reference's `img` set resident on the device, the conv2d
    job of PipelinedConv2dMemFuseTest.cc:249-273 on it) against nb200_conv2d_fused on host images and against the oracle's
    image_flat x kernel_flat block GEMM; twice, so that the cached operand planes are used."""
    rng = np.random.default_rng(7)
    n, c, h, w, kout, k = 3, 5, 20, 17, 6, 3
    imgs = rng.uniform(-0.5, 0.5, size=(n, c, h, w))
    kern = rng.uniform(-0.5, 0.5, size=(kout, c, k, k))
    bias = rng.uniform(-0.5, 0.5, size=kout)
    ref = oracle.transpose_mult_agg(oracle.image_flat(imgs, k, 1, 1, 50, 50), oracle.kernel_flat(kern, bias, 50, 50), 2).dense()
    iset = nb.conv2d_images_create(imgs, "convset", "img")
    for prec, tol in (("bf16x3", 1e-4), ("bf16", 2e-2)):
        host_call = nb.conv2d_fused(imgs, kern, bias, 1, False, nb.PRECISIONS[prec], 50, 50, "convset", "r_host").download_dense()
        for _ in range(2):
            got = nb.conv2d_fused_from_set(iset, kern, bias, 1, False, nb.PRECISIONS[prec], 50, 50, "convset", "r_set").download_dense()
            assert np.array_equal(got, host_call)
            assert np.abs(got - ref).max() <= tol * np.abs(ref).max()
        assert nb.last_kernel_ms() > 0
    iset.destroy()
