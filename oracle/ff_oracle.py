"""ctypes wrapper around oracle/libff_oracle.so plus a dense numpy fp64 cross-check.

TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; never by the product package (netsdb_b200/).

PARITY: pinned against the reference's own UDF headers, compiled unmodified around stand-ins for the object model and
Eigen (oracle/ref/ -> oracle/_ref/libff_ref.so, bound by oracle/ff_ref.py; tests/test_oracle_pinned.py and the vectors
of tests/golden/ref_pinned.npz), not against a reference binary: the reference ships no golden vectors for this path
and its build cannot run here.  The C oracle restates the UDF arithmetic (each function cites file:line in
ff_oracle.h); this module also cross-checks it against a dense numpy evaluation of
softmax(Wo . relu(W1 . X^T + b1) + bo)^T.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libff_oracle.so")


class _Block(C.Structure):
    _fields_ = [("block_row", C.c_int), ("block_col", C.c_int), ("rows", C.c_int), ("cols", C.c_int),
                ("total_rows", C.c_int), ("total_cols", C.c_int), ("data", C.POINTER(C.c_double))]


def build(force: bool = False) -> str:
    """Compile the C oracle (building the checker is not using it)."""
    src = os.path.join(_HERE, "ff_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, i, d = C.c_void_p, C.c_int, C.POINTER(C.c_double)
        L.nbo_synth_value.restype = C.c_double
        L.nbo_synth_value.argtypes = [C.c_uint64, C.c_uint64]
        L.nbo_synth_dense.argtypes = [d, C.c_int64, C.c_int64, C.c_uint64]
        L.nbo_set_new.restype = vp
        L.nbo_set_free.argtypes = [vp]
        L.nbo_set_size.argtypes = [vp]
        L.nbo_set_get.restype = C.POINTER(_Block)
        L.nbo_set_get.argtypes = [vp, i]
        L.nbo_load_matrix.restype = vp
        L.nbo_load_matrix.argtypes = [d, i, i, i, i, i, i]
        L.nbo_load_matrix_synth.restype = vp
        L.nbo_load_matrix_synth.argtypes = [C.c_uint64, i, i, i, i]
        L.nbo_set_to_dense.argtypes = [vp, d, i, i]
        for name in ("nbo_transpose_mult_agg", "nbo_input_layer_join_agg", "nbo_la_transpose_multiply_agg"):
            getattr(L, name).restype = vp
            getattr(L, name).argtypes = [vp, vp, i]
        for name in ("nbo_relu_bias_sum", "nbo_output_layer", "nbo_transpose_bias_sum_as_written", "nbo_output_layer_as_written"):
            getattr(L, name).restype = vp
            getattr(L, name).argtypes = [vp, vp]
        L.nbo_transpose_bias_sum.restype = vp
        L.nbo_transpose_bias_sum.argtypes = [vp, vp, i]
        L.nbo_row_aggregate.restype = vp
        L.nbo_row_aggregate.argtypes = [vp]
        L.nbo_udf_dense2.restype = vp
        L.nbo_udf_dense2.argtypes = [vp, i, i, i, i, d, d, d, d, i]
        L.nbo_lstm_sum.restype = vp
        L.nbo_lstm_sum.argtypes = [vp, vp, vp, i]
        L.nbo_lstm_hidden_state_agg.restype = vp
        L.nbo_lstm_hidden_state_agg.argtypes = [vp, vp, i]
        L.nbo_ff_inference_unit.restype = vp
        L.nbo_ff_inference_unit.argtypes = [vp, vp, vp, vp, vp, i, C.POINTER(vp), C.POINTER(C.c_int32)]
        L.nbo_image_flat.restype = vp
        L.nbo_image_flat.argtypes = [d, i, i, i, i, i, i, i, i, i]
        L.nbo_kernel_flat.restype = vp
        L.nbo_kernel_flat.argtypes = [d, d, i, i, i, i, i]
        L.nbo_conv_result_to_images.argtypes = [vp, i, i, i, i, d]
        L.nbo_dgemm.argtypes = [i, i, i, d, i, d, i, i, d, i]
        L.nbo_hw_threads.restype = i
        _lib = L
    return _lib


def _dptr(a: np.ndarray):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_double))


class BlockSet:
    """Owning handle on an nbo_set (a user set of FFMatrixBlocks)."""

    def __init__(self, ptr):
        if not ptr:
            raise RuntimeError("oracle op failed (block dimension mismatch)")
        self.ptr = C.c_void_p(ptr)

    def __del__(self):
        try:
            if self.ptr:
                lib().nbo_set_free(self.ptr)
                self.ptr = None
        except Exception:
            pass

    def __len__(self):
        return lib().nbo_set_size(self.ptr)

    def blocks(self):
        """Yield (block_row, block_col, total_rows, total_cols, ndarray view rows x cols)."""
        for k in range(len(self)):
            b = lib().nbo_set_get(self.ptr, k).contents
            arr = np.ctypeslib.as_array(b.data, shape=(b.rows, b.cols))
            yield b.block_row, b.block_col, b.total_rows, b.total_cols, arr

    def totals(self):
        b = lib().nbo_set_get(self.ptr, 0).contents
        return b.total_rows, b.total_cols

    def dense(self) -> np.ndarray:
        tr, tc = self.totals()
        out = np.zeros((tr, tc), dtype=np.float64)
        lib().nbo_set_to_dense(self.ptr, _dptr(out), tr, tc)
        return out


def synth_dense(rows: int, cols: int, seed: int) -> np.ndarray:
    out = np.empty((rows, cols), dtype=np.float64)
    lib().nbo_synth_dense(_dptr(out), rows, cols, seed)
    return out


def load_matrix(dense: np.ndarray, block_rows: int, block_cols: int, dont_pad_rows=False, dont_pad_cols=False) -> BlockSet:
    dense = np.ascontiguousarray(dense, dtype=np.float64)
    return BlockSet(lib().nbo_load_matrix(_dptr(dense), dense.shape[0], dense.shape[1], block_rows, block_cols,
                                          int(dont_pad_rows), int(dont_pad_cols)))


def load_matrix_synth(seed, total_rows, total_cols, block_rows, block_cols) -> BlockSet:
    return BlockSet(lib().nbo_load_matrix_synth(seed, total_rows, total_cols, block_rows, block_cols))


def transpose_mult_agg(a: BlockSet, b: BlockSet, threads=1) -> BlockSet:
    return BlockSet(lib().nbo_transpose_mult_agg(a.ptr, b.ptr, threads))


def input_layer_join_agg(a: BlockSet, b: BlockSet, threads=1) -> BlockSet:
    return BlockSet(lib().nbo_input_layer_join_agg(a.ptr, b.ptr, threads))


def la_transpose_multiply_agg(a: BlockSet, b: BlockSet, threads=1) -> BlockSet:
    """LASillyTransposeMultiply1Join + LASillyMultiply2Aggregate: A^T . B."""
    return BlockSet(lib().nbo_la_transpose_multiply_agg(a.ptr, b.ptr, threads))


def relu_bias_sum(a: BlockSet, bias: BlockSet) -> BlockSet:
    return BlockSet(lib().nbo_relu_bias_sum(a.ptr, bias.ptr))


def udf_dense2(a: BlockSet, features, batch, w0, b0, w1, b1, with_totals=True) -> BlockSet:
    """FullyConnectedNetwork / SemanticClassifier: the two-layer network those SelectionComps run on every block."""
    w0 = np.ascontiguousarray(w0, dtype=np.float64); b0 = np.ascontiguousarray(b0, dtype=np.float64)
    w1 = np.ascontiguousarray(w1, dtype=np.float64); b1 = np.ascontiguousarray(b1, dtype=np.float64)
    hidden, labels = w0.shape[0], w1.shape[0]
    assert w0.shape == (hidden, features) and w1.shape == (labels, hidden) and b0.size == hidden and b1.size == labels
    dp = lambda x: x.ctypes.data_as(C.POINTER(C.c_double))
    p = lib().nbo_udf_dense2(a.ptr, features, batch, hidden, labels, dp(w0), dp(b0), dp(w1), dp(b1), int(with_totals))
    if not p:
        raise ValueError("udf_dense2: a block is smaller than features x batch")
    return BlockSet(p)


def transpose_bias_sum(a: BlockSet, bias: BlockSet, apply_exp=True) -> BlockSet:
    return BlockSet(lib().nbo_transpose_bias_sum(a.ptr, bias.ptr, int(apply_exp)))


def transpose_bias_sum_as_written(a: BlockSet, bias: BlockSet) -> BlockSet:
    """FFTransposeBiasSum as this fork writes it (Q1); pinning / documentation only."""
    return BlockSet(lib().nbo_transpose_bias_sum_as_written(a.ptr, bias.ptr))


def output_layer_as_written(a: BlockSet, rowsum: BlockSet) -> BlockSet:
    """FFOutputLayer as this fork writes it (Q2); pinning / documentation only."""
    return BlockSet(lib().nbo_output_layer_as_written(a.ptr, rowsum.ptr))


def row_aggregate(a: BlockSet) -> BlockSet:
    return BlockSet(lib().nbo_row_aggregate(a.ptr))


def output_layer(a: BlockSet, rowsum: BlockSet) -> BlockSet:
    return BlockSet(lib().nbo_output_layer(a.ptr, rowsum.ptr))


def ff_inference_unit(w1, wo, inputs, b1, bo, threads=1):
    """ff::inference_unit.  Returns (output BlockSet, logits BlockSet, labels int32[batch])."""
    batch = inputs.totals()[0]
    labels = np.full(batch, -1, dtype=np.int32)
    logits = C.c_void_p()
    out = lib().nbo_ff_inference_unit(w1.ptr, wo.ptr, inputs.ptr, b1.ptr, bo.ptr, threads, C.byref(logits),
                                      labels.ctypes.data_as(C.POINTER(C.c_int32)))
    return BlockSet(out), BlockSet(logits.value), labels


SIGMOID, TANH = 1, 2  # enum SumActivation (LSTMThreeWaySum.h:12)


def lstm_sum(a: BlockSet, b: BlockSet = None, c: BlockSet = None, activation: int = 0) -> BlockSet:
    """LSTMThreeWaySum / LSTMTwoSum / activation map."""
    return BlockSet(lib().nbo_lstm_sum(a.ptr, b.ptr if b is not None else None, c.ptr if c is not None else None, activation))


def lstm_hidden_state_agg(o_t: BlockSet, c_t: BlockSet, threads=1) -> BlockSet:
    return BlockSet(lib().nbo_lstm_hidden_state_agg(o_t.ptr, c_t.ptr, threads))


def lstm_cell(p: dict, threads=1) -> dict:
    """One LSTM step as src/tests/source/LSTMTest.cc wires it (:204-543).  p maps the reference's set names
    (w_f u_f b_f w_i u_i b_i w_o u_o b_o w_c u_c b_c x_t h_t_1 c_t_1) to BlockSets; returns the sets the driver
    writes: f_t i_t o_t c_t_temp c_t h_t.  Note that the driver combines gates and states with FFInputLayerJoin,
    i.e. as matrix products (:470-510), not element-wise."""
    def gate(w, u, b, act):
        wx = input_layer_join_agg(p[w], p["x_t"], threads)        # :213-225
        uh = input_layer_join_agg(p[u], p["h_t_1"], threads)      # :227-238
        return lstm_sum(wx, uh, p[b], act)                         # :240-247
    r = {}
    r["f_t"] = gate("w_f", "u_f", "b_f", SIGMOID)
    r["i_t"] = gate("w_i", "u_i", "b_i", SIGMOID)                 # :278-312
    r["o_t"] = gate("w_o", "u_o", "b_o", SIGMOID)                 # :343-377
    r["c_t_temp"] = gate("w_c", "u_c", "b_c", TANH)               # :408-440
    fc = input_layer_join_agg(r["f_t"], p["c_t_1"], threads)      # :470-480
    ic = input_layer_join_agg(r["i_t"], r["c_t_temp"], threads)   # :482-494
    r["c_t"] = lstm_sum(fc, ic)                                   # :496-498 LSTMTwoSum
    r["h_t"] = lstm_hidden_state_agg(r["o_t"], r["c_t"], threads)  # :528-543
    return r


def lstm_cell_numpy(d: dict, block: int = 0) -> dict:
    """Dense fp64 evaluation of the same step (independent cross-check).

    block = 0: zero-padding semantics -- a join of a (I x Ka) with b (Kb x J) keeps the matching block
    columns/rows only, which with zero padding is the product over k < min(Ka, Kb).  This is what the reference
    computes whenever every extent is a multiple of the block size.
    block > 0: the reference as written for ragged extents -- every set is carried padded to whole blocks and
    LSTMThreeWaySum applies its activation to the padding too (LSTMThreeWaySum.h:76-95 loops over I*J of the padded
    block), so sigmoid(0) = 0.5 appears in the padding of f_t, i_t, o_t and leaks into c_t and h_t through the
    products (quirk Q9)."""
    def pad(a):
        if not block:
            return a
        r, c = (-a.shape[0]) % block, (-a.shape[1]) % block
        return np.pad(a, ((0, r), (0, c)))
    def mm(a, b):
        k = min(a.shape[1], b.shape[0])
        return a[:, :k] @ b[:k, :]
    sig = lambda s: 1.0 / (1.0 + np.exp(-s))
    e = {k: pad(v) for k, v in d.items()}
    r = {}
    for g, act in (("f", sig), ("i", sig), ("o", sig), ("c", np.tanh)):
        s = mm(e["w_" + g], e["x_t"]) + mm(e["u_" + g], e["h_t_1"]) + e["b_" + g]
        r[{"f": "f_t", "i": "i_t", "o": "o_t", "c": "c_t_temp"}[g]] = act(s)
    r["c_t"] = mm(r["f_t"], e["c_t_1"]) + mm(r["i_t"], r["c_t_temp"])
    r["h_t"] = mm(r["o_t"], np.tanh(r["c_t"]))
    L, B = d["b_f"].shape
    return {k: v[:L, :B] for k, v in r.items()}


def image_flat(images: np.ndarray, k: int, stride: int, padding: int, block_rows: int, block_cols: int) -> BlockSet:
    images = np.ascontiguousarray(images, dtype=np.float64)
    n, c, h, w = images.shape
    return BlockSet(lib().nbo_image_flat(_dptr(images), n, c, h, w, k, stride, padding, block_rows, block_cols))


def kernel_flat(kernels: np.ndarray, bias: np.ndarray, block_rows: int, block_cols: int) -> BlockSet:
    kernels = np.ascontiguousarray(kernels, dtype=np.float64)
    bias = np.ascontiguousarray(bias, dtype=np.float64)
    kout, c, k, k2 = kernels.shape
    assert k == k2
    return BlockSet(lib().nbo_kernel_flat(_dptr(kernels), _dptr(bias), kout, c, k, block_rows, block_cols))


def conv_result_to_images(result: BlockSet, num_images: int, x: int, y: int, z: int) -> np.ndarray:
    """ConvResultToChunks + ConvChunksToImage: images as an (n, z, x, y) array."""
    out = np.zeros((num_images, z, x, y), dtype=np.float64)
    if lib().nbo_conv_result_to_images(result.ptr, num_images, x, y, z, _dptr(out)) != 0:
        raise RuntimeError("conv_result_to_images: total rows not divisible by the number of images, or shape mismatch")
    return out


def dgemm(a: np.ndarray, b: np.ndarray, b_trans: bool) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    m, k = a.shape
    n = b.shape[0] if b_trans else b.shape[1]
    c = np.empty((m, n), dtype=np.float64)
    lib().nbo_dgemm(m, n, k, _dptr(a), a.shape[1], _dptr(b), b.shape[1], int(b_trans), _dptr(c), n)
    return c


def hw_threads() -> int:
    return lib().nbo_hw_threads()


# ----------------------------------------------------------------------------------------------
# numpy restatements (independent of the C code) used to cross-check the C oracle
# ----------------------------------------------------------------------------------------------

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return x ^ (x >> np.uint64(31))


def synth_dense_numpy(rows: int, cols: int, seed: int) -> np.ndarray:
    """Same generator as nbo_synth_value, restated in numpy (FFMatrixUtil.cc:229-237 distribution)."""
    idx = np.arange(rows * cols, dtype=np.uint64)
    with np.errstate(over="ignore"):
        s = _splitmix64(np.array([seed], dtype=np.uint64))[0]
        h = _splitmix64(s ^ (idx * np.uint64(0xD1342543DE82EF95) + np.uint64(1)))
    u = (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    mag = u * (0.5 - 0.0001)
    mag = mag + 0.0001
    return np.where((h & np.uint64(1)) != 0, -mag, mag).reshape(rows, cols)


def ff_dense_numpy(x, w1, b1, wo, bo):
    """softmax(Wo . relu(W1 . X^T + b1) + bo)^T in fp64; returns (probs, logits, labels).

    Mirrors the pipeline of SimpleFF.cc:331-424 without blocking.  probs follows the reference's
    un-stabilised exp(z)/sum(exp(z)): non-finite where exp overflows."""
    y1 = np.maximum(w1 @ x.T + b1.reshape(-1, 1), 0.0)          # hidden x batch
    z = (wo @ y1 + bo.reshape(-1, 1)).T                          # batch x labels
    with np.errstate(over="ignore", invalid="ignore"):
        e = np.exp(z)
        probs = e / e.sum(axis=1, keepdims=True)
    return probs, z, np.argmax(z, axis=1).astype(np.int32)


def im2col_numpy(images: np.ndarray, k: int, stride: int, padding: int) -> np.ndarray:
    """image_flat without blocking: rows img*windows+win, cols ch*k*k+i*k+j, trailing 1.0."""
    n, c, h, w = images.shape
    if padding > 0:
        images = np.pad(images, ((0, 0), (0, 0), (padding, padding), (padding, padding)))
    hp, wp = images.shape[2], images.shape[3]
    vs, hs = (hp - k) // stride + 1, (wp - k) // stride + 1
    out = np.empty((n * vs * hs, c * k * k + 1), dtype=np.float64)
    out[:, -1] = 1.0
    for ch in range(c):
        for i in range(k):
            for j in range(k):
                patch = images[:, ch, i:i + (vs - 1) * stride + 1:stride, j:j + (hs - 1) * stride + 1:stride]
                out[:, ch * k * k + i * k + j] = patch.reshape(-1)
    return out
