"""ctypes binding of libnetsdb_b200.so -- exactly the entry points include/netsdb_b200.h declares.

This is plumbing: no arithmetic happens in Python.  The library must exist (netsdb_b200.build)
and a B200 must be visible; there is no CPU fallback and nothing here imports oracle/.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnetsdb_b200.so")

OK = 0
OP_NT, OP_NN, OP_TN = 0, 1, 2
EPI_NONE, EPI_BIAS_RELU, EPI_BIAS_TRANSPOSE = 0, 1, 2
PREC_BF16, PREC_BF16X3, PREC_FP64, PREC_F16 = 0, 1, 2, 3
PRECISIONS = {"bf16": PREC_BF16, "bf16x3": PREC_BF16X3, "fp64": PREC_FP64, "f16": PREC_F16}

# every symbol include/netsdb_b200.h declares (tests check the .so exports each one)
EXPORTS = [
    "nb200_init", "nb200_shutdown", "nb200_last_error", "nb200_device_count", "nb200_set_stream",
    "nb200_synchronize", "nb200_init_devices", "nb200_use_device", "nb200_current_device", "nb200_launch_count", "nb200_set_option", "nb200_pin_pool", "nb200_unpin_pool", "nb200_set_create",
    "nb200_set_lookup", "nb200_set_destroy", "nb200_set_info", "nb200_upload_block", "nb200_upload_rect",
    "nb200_download_block", "nb200_download_rect", "nb200_set_fill_synth", "nb200_set_invalidate_planes",
    "nb200_set_exp_view", "nb200_set_rename",
    "nb200_gemm_blocks", "nb200_softmax_rows", "nb200_ff_inference_unit", "nb200_conv2d_image_flat",
    "nb200_conv2d_kernel_flat", "nb200_conv2d_fused", "nb200_conv2d_images_create", "nb200_conv2d_fused_from_set", "nb200_transpose_mult_noalloc",
    "nb200_ff_layer1_partial", "nb200_ff_finish_from_layer1", "nb200_set_maxabs", "nb200_sum_blocks", "nb200_udf_dense2", "nb200_conv2d_result_to_images", "nb200_lstm_cell", "nb200_upload_commit", "nb200_set_export_f16", "nb200_set_import_f16",
]


class FFStats(C.Structure):
    _fields_ = [("ms_convert", C.c_float), ("ms_gemm1", C.c_float), ("ms_gemm2", C.c_float),
                ("ms_softmax", C.c_float), ("ms_refine", C.c_float), ("ms_total", C.c_float),
                ("ms_gemm1_kernel", C.c_float), ("refined_rows", C.c_int), ("relevel_rows", C.c_int),
                ("kernel_launches", C.c_int),
                ("splits1", C.c_int)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class LstmSets(C.Structure):
    """nb200_lstm_sets: gate order forget, input, output, cell candidate."""
    _fields_ = [("w", C.c_void_p * 4), ("u", C.c_void_p * 4), ("b", C.c_void_p * 4), ("x_t", C.c_void_p),
                ("h_t_1", C.c_void_p), ("c_t_1", C.c_void_p)]


class Nb200Error(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"nb200 status {status}: {msg}")
        self.status = status


_lib = None


def lib():
    """Load the CUDA library; fail loudly when it is missing (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -m netsdb_b200.build` (nvcc, sm_100a)")
        L = C.CDLL(LIB_PATH)
        vp, i, dp, i64 = C.c_void_p, C.c_int, C.POINTER(C.c_double), C.c_int64
        ip = C.POINTER(C.c_int)
        L.nb200_last_error.restype = C.c_char_p
        L.nb200_init.argtypes = [i]
        L.nb200_device_count.argtypes = [ip]
        L.nb200_set_stream.argtypes = [vp]
        L.nb200_launch_count.restype = i64
        L.nb200_set_option.argtypes = [C.c_char_p, i]
        L.nb200_pin_pool.argtypes = [vp, C.c_size_t]
        L.nb200_unpin_pool.argtypes = [vp]
        L.nb200_set_create.argtypes = [C.c_char_p, C.c_char_p, i, i, i, i, C.POINTER(vp)]
        L.nb200_set_lookup.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(vp)]
        L.nb200_set_destroy.argtypes = [vp]
        L.nb200_set_info.argtypes = [vp, ip, ip, ip, ip]
        L.nb200_upload_block.argtypes = [vp, i, i, vp, i, i]
        L.nb200_upload_rect.argtypes = [vp, i, i, vp, i, i, i64]
        L.nb200_download_block.argtypes = [vp, i, i, vp, i, i]
        L.nb200_download_rect.argtypes = [vp, i, i, vp, i, i, i64]
        L.nb200_set_fill_synth.argtypes = [vp, C.c_uint64, i64]
        L.nb200_set_invalidate_planes.argtypes = [vp]
        L.nb200_set_exp_view.argtypes = [vp, C.c_char_p, C.c_char_p, C.POINTER(vp)]
        L.nb200_set_rename.argtypes = [vp, C.c_char_p, C.c_char_p]
        L.nb200_gemm_blocks.argtypes = [i, vp, vp, i, vp, i, C.c_char_p, C.c_char_p, C.POINTER(vp)]
        L.nb200_softmax_rows.argtypes = [vp, C.c_char_p, C.c_char_p, C.POINTER(vp), vp]
        L.nb200_lstm_cell.argtypes = [C.POINTER(LstmSets), i, C.c_char_p]
        L.nb200_set_export_f16.argtypes = [vp, vp, i64, C.POINTER(C.c_int), vp]
        L.nb200_set_import_f16.argtypes = [vp, vp, i64, i, vp, vp, i]
        L.nb200_sum_blocks.argtypes = [vp, vp, vp, i, C.c_char_p, C.c_char_p, C.POINTER(vp)]
        L.nb200_ff_inference_unit.argtypes = [vp, vp, vp, vp, vp, i, C.c_char_p, C.c_char_p, C.POINTER(vp), vp,
                                              C.POINTER(FFStats)]
        L.nb200_conv2d_image_flat.argtypes = [vp, i, i, i, i, i, i, i, i, i, C.c_char_p, C.c_char_p, C.POINTER(vp)]
        L.nb200_conv2d_kernel_flat.argtypes = [vp, vp, i, i, i, i, i, C.c_char_p, C.c_char_p, C.POINTER(vp)]
        L.nb200_conv2d_fused.argtypes = [vp, i, i, i, i, vp, vp, i, i, i, i, i, i, i, C.c_char_p, C.c_char_p,
                                         C.POINTER(vp)]
        L.nb200_conv2d_result_to_images.argtypes = [vp, i, i, i, i, vp]
        L.nb200_conv2d_images_create.argtypes = [vp, i, i, i, i, C.c_char_p, C.c_char_p, C.POINTER(vp)]
        L.nb200_conv2d_fused_from_set.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, C.c_char_p, C.c_char_p, C.POINTER(vp), C.POINTER(C.c_float)]
        L.nb200_ff_layer1_partial.argtypes = [vp, vp, i, vp, i64]
        L.nb200_ff_finish_from_layer1.argtypes = [vp, i64, i, vp, vp, vp, i, C.c_float, C.c_char_p, C.c_char_p, C.POINTER(vp), vp, ip]
        L.nb200_set_maxabs.argtypes = [vp, C.POINTER(C.c_float)]
        L.nb200_transpose_mult_noalloc.argtypes = [vp, i64, vp, i64, i, i, i, vp, i64, vp]
        _lib = L
    return _lib


def check(status: int):
    if status != OK:
        raise Nb200Error(status, lib().nb200_last_error().decode(errors="replace"))


def init(device: int = 0):
    check(lib().nb200_init(device))


def init_devices(devices):
    arr = (C.c_int * len(devices))(*devices)
    check(lib().nb200_init_devices(arr, len(devices)))


def use_device(device: int):
    """The calling thread works on `device` from now on (thread-local)."""
    check(lib().nb200_use_device(device))


def shutdown():
    check(lib().nb200_shutdown())


def synchronize():
    check(lib().nb200_synchronize())


def set_stream(stream_ptr: int | None):
    check(lib().nb200_set_stream(C.c_void_p(stream_ptr or 0)))


def set_option(key: str, value: int):
    check(lib().nb200_set_option(key.encode(), int(value)))


def launch_count() -> int:
    return int(lib().nb200_launch_count())


def device_count() -> int:
    n = C.c_int(0)
    check(lib().nb200_device_count(C.byref(n)))
    return n.value


def _b(s: str) -> bytes:
    return s.encode()


def _hostptr(a: np.ndarray):
    return C.c_void_p(a.ctypes.data)


class DeviceSet:
    """Handle on a device-resident set of FFMatrixBlocks (nb200_set*)."""

    def __init__(self, ptr):
        self.ptr = C.c_void_p(ptr) if not isinstance(ptr, C.c_void_p) else ptr

    @staticmethod
    def create(db: str, name: str, total_rows: int, total_cols: int, block_rows: int, block_cols: int) -> "DeviceSet":
        p = C.c_void_p()
        check(lib().nb200_set_create(_b(db), _b(name), total_rows, total_cols, block_rows, block_cols, C.byref(p)))
        return DeviceSet(p)

    @staticmethod
    def lookup(db: str, name: str) -> "DeviceSet":
        p = C.c_void_p()
        check(lib().nb200_set_lookup(_b(db), _b(name), C.byref(p)))
        return DeviceSet(p)

    def info(self):
        a, b, c_, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        check(lib().nb200_set_info(self.ptr, C.byref(a), C.byref(b), C.byref(c_), C.byref(d)))
        return a.value, b.value, c_.value, d.value

    def destroy(self):
        check(lib().nb200_set_destroy(self.ptr))
        self.ptr = None

    def upload_block(self, block_row: int, block_col: int, block: np.ndarray):
        assert block.dtype == np.float64 and block.flags["C_CONTIGUOUS"] and block.ndim == 2
        check(lib().nb200_upload_block(self.ptr, block_row, block_col, _hostptr(block), block.shape[0], block.shape[1]))

    def upload_rect(self, row0: int, col0: int, rect: np.ndarray):
        assert rect.dtype == np.float64 and rect.ndim == 2 and rect.strides[1] == 8
        check(lib().nb200_upload_rect(self.ptr, row0, col0, _hostptr(rect), rect.shape[0], rect.shape[1],
                                      rect.strides[0] // 8))

    def f16_ldk(self) -> int:
        return (self.info()[1] + 63) // 64 * 64

    def export_f16(self, plane: np.ndarray | None = None, row_sumsq: np.ndarray | None = None):
        """nb200_set_export_f16: (plane uint16[rows, ldk], exponent, row_sumsq float32[rows]) of a resident fp64 set."""
        rows, ldk = self.info()[0], self.f16_ldk()
        plane = np.empty((rows, ldk), dtype=np.uint16) if plane is None else plane
        row_sumsq = np.empty(rows, dtype=np.float32) if row_sumsq is None else row_sumsq
        e = C.c_int()
        check(lib().nb200_set_export_f16(self.ptr, _hostptr(plane), ldk, C.byref(e), _hostptr(row_sumsq)))
        return plane, e.value, row_sumsq

    def import_f16(self, plane: np.ndarray, exp: int, row_sumsq: np.ndarray, block_ptrs: np.ndarray, edge_ld: int):
        """nb200_set_import_f16: block_ptrs = uint64 array of the addresses of the set's fp64 host blocks (row-major block
        grid, pinned memory).  All arrays must stay alive and unmodified until the next synchronising call."""
        assert plane.dtype == np.uint16 and plane.flags["C_CONTIGUOUS"] and row_sumsq.dtype == np.float32
        assert block_ptrs.dtype == np.uint64 and block_ptrs.flags["C_CONTIGUOUS"]
        self._keep = (plane, row_sumsq, block_ptrs)
        check(lib().nb200_set_import_f16(self.ptr, _hostptr(plane), plane.shape[1], int(exp), _hostptr(row_sumsq),
                                         _hostptr(block_ptrs), int(edge_ld)))

    def upload_commit(self):
        """nb200_upload_commit: readers of this set wait for its uploads only."""
        check(lib().nb200_upload_commit(self.ptr))

    def upload_dense(self, dense: np.ndarray):
        """ff::loadMatrix analogue: send the matrix block by block (full-size zero-padded blocks)."""
        tr, tc, br, bc = self.info()
        assert dense.shape == (tr, tc)
        blk = np.zeros((br, bc), dtype=np.float64)
        for i in range((tr + br - 1) // br):
            for j in range((tc + bc - 1) // bc):
                blk[:] = 0.0
                sub = dense[i * br:(i + 1) * br, j * bc:(j + 1) * bc]
                blk[:sub.shape[0], :sub.shape[1]] = sub
                self.upload_block(i, j, blk)
                synchronize()  # blk is reused

    def download_block(self, block_row: int, block_col: int, rows: int | None = None, cols: int | None = None):
        _, _, br, bc = self.info()
        rows, cols = rows or br, cols or bc
        out = np.empty((rows, cols), dtype=np.float64)
        check(lib().nb200_download_block(self.ptr, block_row, block_col, _hostptr(out), rows, cols))
        return out

    def download_dense(self) -> np.ndarray:
        tr, tc, _, _ = self.info()
        out = np.empty((tr, tc), dtype=np.float64)
        if tr and tc:
            check(lib().nb200_download_rect(self.ptr, 0, 0, _hostptr(out), tr, tc, tc))
        return out

    def fill_synth(self, seed: int, row_offset: int = 0):
        check(lib().nb200_set_fill_synth(self.ptr, seed, row_offset))

    def invalidate_planes(self):
        check(lib().nb200_set_invalidate_planes(self.ptr))


def gemm_blocks(op: int, a: DeviceSet, b: DeviceSet, epilogue: int, bias: DeviceSet | None, precision: int, db: str,
                out_set: str) -> DeviceSet:
    p = C.c_void_p()
    check(lib().nb200_gemm_blocks(op, a.ptr, b.ptr, epilogue, bias.ptr if bias else None, precision, _b(db),
                                  _b(out_set), C.byref(p)))
    return DeviceSet(p)


def conv2d_result_to_images(result: DeviceSet, num_images: int, x: int, y: int, z: int) -> np.ndarray:
    """ConvResultToChunks + ConvChunksToImage: the result matrix re-assembled as an (n, z, x, y) fp64 array."""
    out = np.empty((num_images, z, x, y), dtype=np.float64)
    check(lib().nb200_conv2d_result_to_images(result.ptr, num_images, x, y, z, _hostptr(out)))
    return out


def udf_dense2(a: DeviceSet, features: int, batch: int, w0: np.ndarray, b0: np.ndarray, w1: np.ndarray, b1: np.ndarray,
               db: str = "ff", out_set: str = "output") -> DeviceSet:
    """FullyConnectedNetwork / SemanticClassifier (nb200_udf_dense2): the two-layer network inside one SelectionComp."""
    w0 = np.ascontiguousarray(w0, dtype=np.float64); b0 = np.ascontiguousarray(b0, dtype=np.float64)
    w1 = np.ascontiguousarray(w1, dtype=np.float64); b1 = np.ascontiguousarray(b1, dtype=np.float64)
    hidden, labels = w0.shape[0], w1.shape[0]
    if w0.shape != (hidden, features) or w1.shape != (labels, hidden) or b0.size != hidden or b1.size != labels:
        raise ValueError("udf_dense2: weight shapes do not match (hidden x features, hidden, labels x hidden, labels)")
    p = C.c_void_p()
    check(lib().nb200_udf_dense2(a.ptr, features, batch, hidden, labels, _hostptr(w0), _hostptr(b0), _hostptr(w1), _hostptr(b1),
                                 _b(db), _b(out_set), C.byref(p)))
    return DeviceSet(p)


ACT_NONE, ACT_SIGMOID, ACT_TANH = 0, 1, 2


def sum_blocks(a: DeviceSet, b: DeviceSet | None, c: DeviceSet | None, activation: int, db: str, out_set: str) -> DeviceSet:
    """LSTMThreeWaySum / LSTMTwoSum / activation map (nb200_sum_blocks)."""
    p = C.c_void_p()
    check(lib().nb200_sum_blocks(a.ptr, b.ptr if b else None, c.ptr if c else None, activation, _b(db), _b(out_set),
                                 C.byref(p)))
    return DeviceSet(p)


def lstm_cell(db: str, precision: int) -> dict:
    """nb200_lstm_cell on the reference driver's set names in `db`; returns the six result sets."""
    L = DeviceSet.lookup
    a = LstmSets()
    for k, g in enumerate("fioc"):
        a.w[k], a.u[k], a.b[k] = L(db, "w_" + g).ptr.value, L(db, "u_" + g).ptr.value, L(db, "b_" + g).ptr.value
    a.x_t, a.h_t_1, a.c_t_1 = L(db, "x_t").ptr.value, L(db, "h_t_1").ptr.value, L(db, "c_t_1").ptr.value
    check(lib().nb200_lstm_cell(C.byref(a), precision, _b(db)))
    return {n: L(db, n) for n in ("f_t", "i_t", "o_t", "c_t_temp", "c_t", "h_t")}


def softmax_rows(logits: DeviceSet, db: str, out_set: str, want_labels: bool = True):
    p = C.c_void_p()
    rows = logits.info()[0]
    labels = np.empty(rows, dtype=np.int32) if want_labels else None
    check(lib().nb200_softmax_rows(logits.ptr, _b(db), _b(out_set), C.byref(p),
                                   _hostptr(labels) if want_labels else None))
    return DeviceSet(p), labels


def ff_inference_unit(w1: DeviceSet, wo: DeviceSet, inputs: DeviceSet, b1: DeviceSet, bo: DeviceSet, precision: int,
                      db: str = "ff", out_set: str = "output", want_labels: bool = True):
    """Returns (output DeviceSet, labels int32[batch] or None, stats dict)."""
    p = C.c_void_p()
    rows = inputs.info()[0]
    labels = np.empty(rows, dtype=np.int32) if want_labels else None
    st = FFStats()
    check(lib().nb200_ff_inference_unit(w1.ptr, wo.ptr, inputs.ptr, b1.ptr, bo.ptr, precision, _b(db), _b(out_set),
                                        C.byref(p), _hostptr(labels) if want_labels else None, C.byref(st)))
    return DeviceSet(p), labels, st.as_dict()


def conv2d_image_flat(images: np.ndarray, k: int, stride: int, padding: int, block_rows: int, block_cols: int,
                      db: str = "conv2d", out_set: str = "image_flat") -> DeviceSet:
    images = np.ascontiguousarray(images, dtype=np.float64)
    n, c_, h, w = images.shape
    p = C.c_void_p()
    check(lib().nb200_conv2d_image_flat(_hostptr(images), n, c_, h, w, k, stride, padding, block_rows, block_cols,
                                        _b(db), _b(out_set), C.byref(p)))
    return DeviceSet(p)


def conv2d_kernel_flat(kernels: np.ndarray, bias: np.ndarray, block_rows: int, block_cols: int, db: str = "conv2d",
                       out_set: str = "kernel_flat") -> DeviceSet:
    kernels = np.ascontiguousarray(kernels, dtype=np.float64)
    bias = np.ascontiguousarray(bias, dtype=np.float64)
    kout, c_, k, _ = kernels.shape
    p = C.c_void_p()
    check(lib().nb200_conv2d_kernel_flat(_hostptr(kernels), _hostptr(bias), kout, c_, k, block_rows, block_cols,
                                         _b(db), _b(out_set), C.byref(p)))
    return DeviceSet(p)


def conv2d_fused(images: np.ndarray, kernels: np.ndarray, bias: np.ndarray, padding: int, relu: bool, precision: int,
                 block_rows: int, block_cols: int, db: str = "conv2d", out_set: str = "result") -> DeviceSet:
    images = np.ascontiguousarray(images, dtype=np.float64)
    kernels = np.ascontiguousarray(kernels, dtype=np.float64)
    bias = np.ascontiguousarray(bias, dtype=np.float64)
    n, c_, h, w = images.shape
    kout, _, k, _ = kernels.shape
    p = C.c_void_p()
    check(lib().nb200_conv2d_fused(_hostptr(images), n, c_, h, w, _hostptr(kernels), _hostptr(bias), kout, k, padding,
                                   int(relu), precision, block_rows, block_cols, _b(db), _b(out_set), C.byref(p)))
    return DeviceSet(p)


_last_kernel_ms = 0.0


def conv2d_images_create(images: np.ndarray, db: str = "conv2d", set_name: str = "img") -> DeviceSet:
    images = np.ascontiguousarray(images, dtype=np.float64)
    n, c, h, w = images.shape
    p = C.c_void_p()
    check(lib().nb200_conv2d_images_create(_hostptr(images), n, c, h, w, _b(db), _b(set_name), C.byref(p)))
    return DeviceSet(p)


def conv2d_fused_from_set(images: DeviceSet, kernels: np.ndarray, bias: np.ndarray, padding: int, relu: bool, precision: int,
                          block_rows: int, block_cols: int, db: str = "conv2d", out_set: str = "result") -> DeviceSet:
    global _last_kernel_ms
    kernels = np.ascontiguousarray(kernels, dtype=np.float64)
    bias = np.ascontiguousarray(bias, dtype=np.float64)
    kout, _, k, _ = kernels.shape
    p = C.c_void_p()
    ms = C.c_float(0)
    check(lib().nb200_conv2d_fused_from_set(images.ptr, _hostptr(kernels), _hostptr(bias), kout, k, padding, int(relu), precision,
                                            block_rows, block_cols, _b(db), _b(out_set), C.byref(p), C.byref(ms)))
    _last_kernel_ms = ms.value
    return DeviceSet(p)


def last_kernel_ms() -> float:
    return _last_kernel_ms
