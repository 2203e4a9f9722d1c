"""Host-side mirror of the reference's ff:: driver namespace (FF/source/SimpleFF.cc,
FF/source/FFMatrixUtil.cc) on top of the C ABI.  Same names, argument meaning and error behaviour:
failures raise (the reference prints and exit(1)s).  No arithmetic happens here.
"""
from __future__ import annotations

import numpy as np

from . import capi

_SET_SHAPES: dict[tuple[str, str], tuple[int, int, int, int]] = {}


def createDatabase(db: str) -> None:  # SimpleFF.cc:58-66 -- databases are implicit on the device
    return None


def setup(db: str) -> None:  # SimpleFF.cc:69-82 registers the UDF .so's; nothing to ship here
    return None


def createSet(db: str, set_name: str, set_type: str = "FFMatrixBlock", page_mb: int = 64) -> None:
    """ff::createSet(pdbClient, db, set, name, size) (SimpleFF.cc:29-55).  The device set is sized when
    loadMatrix learns the matrix dimensions, as the reference's set is filled by sendData."""
    _SET_SHAPES.setdefault((db, set_name), (0, 0, 1, 1))


def loadMatrix(db: str, set_name: str, totalX: int, totalY: int, blockX: int, blockY: int, dont_pad_x: bool = False,
               dont_pad_y: bool = False, seed: int = 0, partitionByCol: bool = True, row_offset: int = 0,
               dense: np.ndarray | None = None) -> capi.DeviceSet:
    """ff::loadMatrix (FFMatrixUtil.cc:216-328).  Values come from the seeded device generator that
    reproduces the reference's distribution (sign * U(1e-4, 0.5)), or from `dense` when given, in which
    case the matrix is sent block by block like the reference's sendData loop."""
    if totalX == 0 or totalY == 0 or blockX == 0 or blockY == 0:  # FFMatrixUtil.cc:221-225
        return None
    s = capi.DeviceSet.create(db, set_name, totalX, totalY, blockX, blockY)
    if dense is None:
        s.fill_synth(seed, row_offset)
    else:
        s.upload_dense(np.ascontiguousarray(dense, dtype=np.float64))
    _SET_SHAPES[(db, set_name)] = (totalX, totalY, blockX, blockY)
    return s


def inference_unit(db: str, w1: str, wo: str, inputs: str, b1: str, bo: str, output: str, dropout_rate: float = 0.0,
                   precision: str = "bf16x3"):
    """ff::inference_unit (SimpleFF.cc:331-424) as one fused device stage.

    dropout_rate must be 0: the reference's dropout draws from a time-seeded rand()
    (FFReluBiasSum.h:67,73-80) and cannot be reproduced."""
    if dropout_rate != 0.0:
        raise ValueError("dropout_rate must be 0 (the reference's dropout is time-seeded and irreproducible)")
    L = capi.DeviceSet.lookup
    out, labels, stats = capi.ff_inference_unit(L(db, w1), L(db, wo), L(db, inputs), L(db, b1), L(db, bo),
                                                capi.PRECISIONS[precision], db, output)
    return out, labels, stats
