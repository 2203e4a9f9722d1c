"""K-sharded FF inference across GPUs (the fallback for weight sets larger than one GPU's HBM).

Every rank holds columns [c0, c1) of `inputs` (all batch rows) and of `w1`; b1, wo, bo are replicated.  Layer 1 is a
partial product per rank, summed over ranks with ONE NCCL collective over NVLink (reduce-scatter by rows, so that
each rank continues with batch/world rows: bias + ReLU, layer 2 and the softmax are row-local again).  This mirrors
the reference's hash partitioning by blockColIndex (FFMatrixBlock.h:151-156) and its cross-worker aggregation
(PipelineStage.cc:120-205), with the TCP shuffle replaced by the collective.

torch.distributed is plumbing only: device buffers for the partials and the collective.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .shard import row_shard


def col_shard(total_cols: int, block_cols: int, world: int, rank: int) -> tuple[int, int]:
    """(first column, column count) of `rank`, cut at block-column granularity like the reference's partitions."""
    nblk = (total_cols + block_cols - 1) // block_cols
    b0, nb = row_shard(nblk, world, rank)
    c0 = b0 * block_cols
    return c0, max(0, min(total_cols, (b0 + nb) * block_cols) - c0)


def ff_inference_k_sharded(dist, torch, rank: int, world: int, w1_shard, inputs_shard, b1, wo, bo, precision: int,
                           db: str = "ff", out_set: str = "output", timing: dict | None = None):
    """Returns (output DeviceSet holding this rank's rows, labels of this rank's rows, undecided row count, row0).
    `timing` (optional) receives the device time of the collective in ms (CUDA events on the current stream) and its payload."""
    B = inputs_shard.info()[0]
    H = w1_shard.info()[0]
    Bp = (B + world - 1) // world * world                      # reduce-scatter wants equal chunks: pad the batch
    partial = torch.zeros((Bp, H + 1), dtype=torch.float32, device="cuda")
    capi.check(capi.lib().nb200_ff_layer1_partial(w1_shard.ptr, inputs_shard.ptr, precision, C.c_void_p(partial.data_ptr()), H + 1))
    mx = C.c_float(0)
    capi.check(capi.lib().nb200_set_maxabs(w1_shard.ptr, C.byref(mx)))
    capi.synchronize()
    torch.cuda.synchronize()
    rows_per = Bp // world
    mine = torch.empty((rows_per, H + 1), dtype=torch.float32, device="cuda")
    wmax = torch.tensor([mx.value], dtype=torch.float32, device="cuda")
    ev = None
    if timing is not None and torch.cuda.is_available():
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record()
    if world > 1:
        if dist.get_backend() == "nccl":
            dist.reduce_scatter_tensor(mine, partial, op=dist.ReduceOp.SUM)     # the one data-path collective
        else:                                                                   # gloo (CPU tests): all-reduce + slice
            dist.all_reduce(partial, op=dist.ReduceOp.SUM)
            mine.copy_(partial[rank * rows_per:(rank + 1) * rows_per])
        dist.all_reduce(wmax, op=dist.ReduceOp.MAX)
    else:
        mine.copy_(partial)
    if ev is not None:
        ev[1].record()
    torch.cuda.synchronize()
    if ev is not None:
        timing["collective_ms"] = ev[0].elapsed_time(ev[1])
        timing["payload_bytes"] = int(partial.numel() * 4)
    row0 = rank * rows_per
    rows = max(0, min(B, row0 + rows_per) - row0)
    labels = np.empty(rows, dtype=np.int32)
    out = C.c_void_p()
    und = C.c_int(0)
    capi.check(capi.lib().nb200_ff_finish_from_layer1(C.c_void_p(mine.data_ptr()), H + 1, rows, b1.ptr, wo.ptr, bo.ptr, precision,
                                                      float(wmax.item()), db.encode(), out_set.encode(), C.byref(out),
                                                      C.c_void_p(labels.ctypes.data), C.byref(und)))
    return capi.DeviceSet(out), labels, und.value, row0
