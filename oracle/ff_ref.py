"""ctypes binding of oracle/_ref/libff_ref.so: the reference's OWN UDF headers (netsDB src/FF, src/LSTM,
src/conv2d_memory_fusion, src/sharedLibraries), compiled unmodified by `make -C oracle ref` around stand-ins for the
object model and Eigen (oracle/ref/).  TEST INFRASTRUCTURE: it pins the restatement in ff_oracle.c and generates the
golden vectors under tests/golden/; nothing in the product may load it.

The library exists only where the reference tree was mounted when it was built (it travels with the repository
snapshot but is not in git); `available()` says whether it is there.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
import subprocess
import sys

import numpy as np

from . import ff_oracle as fo

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "libff_ref.so")
_lib = None


def build() -> bool:
    """Compile the reference's UDF headers if the tree is mounted; returns whether the library exists afterwards."""
    if os.path.isdir("/root/reference/src/FF/headers"):
        fo.build()
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])
    return os.path.exists(_LIB_PATH)


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        fo.lib()  # libff_ref.so links against libff_oracle.so for the nbo_set containers
        L = C.CDLL(_LIB_PATH)
        vp, i, d = C.c_void_p, C.c_int, C.POINTER(C.c_double)
        L.ref_join_agg.restype = vp
        L.ref_join_agg.argtypes = [i, vp, vp]
        L.ref_bias_join.restype = vp
        L.ref_bias_join.argtypes = [i, vp, vp, C.c_double]
        L.ref_row_aggregate.restype = vp
        L.ref_row_aggregate.argtypes = [vp]
        L.ref_lstm_sum.restype = vp
        L.ref_lstm_sum.argtypes = [vp, vp, vp, i]
        L.ref_la_multiply_agg.restype = vp
        L.ref_la_multiply_agg.argtypes = [i, vp, vp]
        L.ref_image_flat.restype = vp
        L.ref_image_flat.argtypes = [d, i, i, i, i, i, i, i, i, i]
        L.ref_kernel_flat.restype = vp
        L.ref_kernel_flat.argtypes = [d, vp, i, i, i, i, i, i]
        L.ref_conv_result_to_images.argtypes = [vp, i, i, i, i, d]
        L.ref_fully_connected.restype = vp
        L.ref_fully_connected.argtypes = [vp, i, i, i, i, d, d, d, d]
        L.ref_semantic_classifier.restype = vp
        L.ref_semantic_classifier.argtypes = [vp, i, i, i]
        _lib = L
    return _lib


@contextlib.contextmanager
def _quiet():
    """The reference's UDFs print per element in this fork (FFTransposeMult.h:94-99 and others)."""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    try:
        yield
    finally:
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)


def _set(ptr) -> fo.BlockSet:
    return fo.BlockSet(ptr)


def _d(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def transpose_mult_agg(a, b):       # FFTransposeMult + FFAggMatrix
    with _quiet():
        return _set(lib().ref_join_agg(0, a.ptr, b.ptr))


def input_layer_join_agg(a, b):     # FFInputLayerJoin + FFAggMatrix
    with _quiet():
        return _set(lib().ref_join_agg(1, a.ptr, b.ptr))


def lstm_hidden_state_agg(o, c):    # LSTMHiddenState + FFAggMatrix (note: tanh'es c's blocks in place, once per pair)
    with _quiet():
        return _set(lib().ref_join_agg(2, o.ptr, c.ptr))


def relu_bias_sum(a, bias, dropout_rate=0.0):
    with _quiet():
        return _set(lib().ref_bias_join(0, a.ptr, bias.ptr, dropout_rate))


def transpose_bias_sum_as_written(a, bias):
    with _quiet():
        return _set(lib().ref_bias_join(1, a.ptr, bias.ptr, 0.0))


def output_layer_as_written(a, rowsum):
    with _quiet():
        return _set(lib().ref_bias_join(2, a.ptr, rowsum.ptr, 0.0))


def row_aggregate(a):
    with _quiet():
        return _set(lib().ref_row_aggregate(a.ptr))


def lstm_sum(a, b, c=None, activation=0):
    with _quiet():
        return _set(lib().ref_lstm_sum(a.ptr, b.ptr, c.ptr if c is not None else None, activation))


def la_multiply_agg(a, b, transpose_left=False):
    with _quiet():
        return _set(lib().ref_la_multiply_agg(int(transpose_left), a.ptr, b.ptr))


def image_flat(images: np.ndarray, k: int, stride: int, padding: int, block_x: int, block_y: int):
    images = np.ascontiguousarray(images, dtype=np.float64)
    n, c, h, w = images.shape
    with _quiet():
        return _set(lib().ref_image_flat(_d(images), n, c, h, w, k, stride, padding, block_x, block_y))


def kernel_flat(kernels: np.ndarray, bias_set, block_x: int, block_y: int, with_join: bool = True):
    kernels = np.ascontiguousarray(kernels, dtype=np.float64)
    kout, c, k, _ = kernels.shape
    with _quiet():
        return _set(lib().ref_kernel_flat(_d(kernels), bias_set.ptr if bias_set is not None else None, kout, c, k, block_x, block_y,
                                          int(with_join)))


def conv_result_to_images(result, n: int, x: int, y: int, z: int) -> np.ndarray:
    out = np.zeros((n, z, x, y), dtype=np.float64)
    with _quiet():
        got = lib().ref_conv_result_to_images(result.ptr, n, x, y, z, _d(out))
    if got < 0:
        raise RuntimeError("conv_result_to_images: image index out of range")
    return out


def fully_connected(a, features: int, labels: int, hidden: int, batch: int):
    """FullyConnectedNetwork constructed as FCProjTest.cc does; returns (output set, w0, b0, w1, b1) -- the weights the
    reference drew from std::random_device for this run."""
    w0, b0 = np.empty((hidden, features)), np.empty(hidden)
    w1, b1 = np.empty((labels, hidden)), np.empty(labels)
    with _quiet():
        out = _set(lib().ref_fully_connected(a.ptr, features, labels, hidden, batch, _d(w0), _d(b0), _d(w1), _d(b1)))
    return out, w0, b0, w1, b1


def semantic_classifier(a, size_embed: int, dense0: int, dense1: int):
    with _quiet():
        return _set(lib().ref_semantic_classifier(a.ptr, size_embed, dense0, dense1))
