"""CPU tests of the multi-GPU host logic: row sharding, block rectangles, and -- with two gloo processes --
that sharding the batch by rows with replicated weights reproduces the unsharded pipeline exactly (the FF
path is row-local, so no data-path collective is needed) and that labels gather in global row order."""
import os
import socket
import sys

import numpy as np
import pytest

from netsdb_b200.shard import block_rects, gather_labels, row_shard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_row_shard_covers_every_row_once():
    for total in (0, 1, 7, 1000, 10000):
        for world in (1, 2, 3, 4, 8):
            cuts = [row_shard(total, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and sum(n for _, n in cuts) == total
            for (a0, an), (b0, _) in zip(cuts, cuts[1:]):
                assert a0 + an == b0
            assert max(n for _, n in cuts) - min(n for _, n in cuts) <= 1
    assert row_shard(10000, 8, 3) == (3750, 1250)   # BASELINE config 3: 1,250 rows per GPU at 8-way


def test_block_rects_tile_the_shard():
    total_r, total_c, br, bc = 2300, 1700, 1000, 600
    dense = np.arange(total_r * total_c, dtype=np.float64).reshape(total_r, total_c)
    for world in (1, 2, 3, 8):
        seen = np.zeros(total_r, dtype=np.int64)
        for rank in range(world):
            row0, nrows = row_shard(total_r, world, rank)
            shard = np.full((nrows, total_c), np.nan)
            for b_r, b_c, r_in, rows, s_row, col0, cols in block_rects(total_r, total_c, br, bc, world, rank):
                block = np.zeros((br, bc))
                sub = dense[b_r * br:(b_r + 1) * br, b_c * bc:(b_c + 1) * bc]
                block[:sub.shape[0], :sub.shape[1]] = sub                      # a reference-style padded host block
                shard[s_row:s_row + rows, col0:col0 + cols] = block[r_in:r_in + rows, :cols]
                if b_c == 0:
                    seen[row0 + s_row: row0 + s_row + rows] += 1
            assert np.array_equal(shard, dense[row0:row0 + nrows])
        assert (seen == 1).all()


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from oracle import ff_oracle as fo

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    batch, feat, hid, lab, blk = 203, 150, 60, 9, 50
    x = fo.synth_dense(batch, feat, 1)
    w1, b1 = fo.synth_dense(hid, feat, 2), fo.synth_dense(hid, 1, 3)
    wo, bo = fo.synth_dense(lab, hid, 4), fo.synth_dense(lab, 1, 5)
    row0, nrows = row_shard(batch, world, rank)
    # each rank assembles ITS rows from the reference-style blocks of the full matrix
    shard = np.zeros((nrows, feat))
    full_blocks = {(br, bc): arr.copy() for br, bc, _, _, arr in fo.load_matrix(x, blk, blk).blocks()}
    for b_r, b_c, r_in, rows, s_row, col0, cols in block_rects(batch, feat, blk, blk, world, rank):
        shard[s_row:s_row + rows, col0:col0 + cols] = full_blocks[(b_r, b_c)][r_in:r_in + rows, :cols]
    L = lambda d, bc_: fo.load_matrix(d, blk, bc_)
    _, logits, labels = fo.ff_inference_unit(L(w1, blk), L(wo, blk), L(shard, blk), L(b1, 1), L(bo, 1), threads=1)
    all_labels = gather_labels(labels, batch, world, rank, dist)
    ok = True
    if rank == 0:
        _, ref_logits, ref_labels = fo.ff_inference_unit(L(w1, blk), L(wo, blk), L(x, blk), L(b1, 1), L(bo, 1), threads=1)
        ok = bool(np.array_equal(all_labels, ref_labels)) and bool(
            np.abs(logits.dense() - ref_logits.dense()[row0:row0 + nrows]).max() <= 1e-12)
    q.put((rank, ok))
    dist.barrier()
    dist.destroy_process_group()


def test_row_sharded_pipeline_equals_unsharded_world2_gloo():
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert results == {0: True, 1: True}


def test_col_shard_is_block_aligned_and_complete():
    from netsdb_b200.ksharded import col_shard

    for total, blk in ((597540, 1000), (9000, 500), (1024, 100), (7, 3)):
        for world in (1, 2, 4, 8):
            cuts = [col_shard(total, blk, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and sum(n for _, n in cuts) == total
            for (a0, an), (b0, _) in zip(cuts, cuts[1:]):
                assert a0 + an == b0 and b0 % blk == 0 or b0 >= total
