"""netsdb_b200 -- B200 (sm_100a) implementation of netsDB's block-partitioned FF inference path.

The product is libnetsdb_b200.so (hand-written CUDA behind the C ABI of include/netsdb_b200.h);
this package is its Python host-side mirror: `capi` binds the C ABI with ctypes, `ff` mirrors the
reference's ff:: driver functions (FF/headers/SimpleFF.h:14-52, FFMatrixUtil.h:14-40) by name.
"""
from . import capi  # noqa: F401

__all__ = ["capi", "ff", "build"]
