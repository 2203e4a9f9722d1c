"""Row sharding of the FF path across GPUs.

Output block (i, j) of FFTransposeMult depends only on row block i of one operand and row block j of the
other, and every later stage is row-local, so the batch rows of `inputs` split contiguously across ranks
with w1, b1, wo, bo replicated and NO collective on the data path.  Shards are cut at row granularity (not
block granularity: 10 row blocks do not divide by 4 or 8), so a reference block can straddle two ranks;
`block_rects` lists the rectangles of each host block a rank uploads with nb200_upload_rect.

The only communication is control-plane: gathering the per-rank labels for the client, and the barrier /
max-over-ranks reduction of bench.py.
"""
from __future__ import annotations


def row_shard(total_rows: int, world: int, rank: int) -> tuple[int, int]:
    """(first row, row count) of `rank`: contiguous, sizes differ by at most one, earlier ranks get the extra row."""
    base, extra = divmod(total_rows, world)
    rows = base + (1 if rank < extra else 0)
    row0 = rank * base + min(rank, extra)
    return row0, rows


def block_rects(total_rows: int, total_cols: int, block_rows: int, block_cols: int, world: int, rank: int):
    """Yield (block_row, block_col, r0_in_block, rows, shard_row0, col0, cols) for every piece of a host block that
    belongs to `rank`: rows [r0_in_block, r0_in_block + rows) of block (block_row, block_col) land at rows
    [shard_row0, ...) of the rank's device set, columns [col0, col0 + cols)."""
    row0, nrows = row_shard(total_rows, world, rank)
    if nrows == 0:
        return
    first_br, last_br = row0 // block_rows, (row0 + nrows - 1) // block_rows
    ncb = (total_cols + block_cols - 1) // block_cols
    for br in range(first_br, last_br + 1):
        lo = max(row0, br * block_rows)
        hi = min(row0 + nrows, (br + 1) * block_rows, total_rows)
        for bc in range(ncb):
            cols = min(block_cols, total_cols - bc * block_cols)
            yield br, bc, lo - br * block_rows, hi - lo, lo - row0, bc * block_cols, cols


def gather_labels(local_labels, total_rows: int, world: int, rank: int, dist=None):
    """All ranks' labels in global row order (a control-plane gather; `dist` = torch.distributed or None for world 1)."""
    import numpy as np

    if world == 1 or dist is None:
        return np.asarray(local_labels, dtype=np.int32)
    import torch

    sizes = [row_shard(total_rows, world, r)[1] for r in range(world)]
    mx = max(sizes)
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"   # NCCL moves device tensors only
    buf = torch.full((mx,), -1, dtype=torch.int32, device=dev)
    buf[: sizes[rank]] = torch.as_tensor(np.asarray(local_labels, dtype=np.int32)).to(dev)
    out = [torch.empty(mx, dtype=torch.int32, device=dev) for _ in range(world)]
    dist.all_gather(out, buf)
    return np.concatenate([o[:n].cpu().numpy() for o, n in zip(out, sizes)])
