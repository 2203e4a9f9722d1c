"""Full-size GPU tests (BASELINE.json configs[1] and the word2vec shape): the oracle cannot evaluate 1.2e13 flop in a
test, so parity is checked through size-independent properties and on a SAMPLE of rows recomputed in fp64 on the host
from the same seeded generator: label equality, logit tolerance, probabilities summing to one, idempotence."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rows(oracle, rows, cols, seed, idx):
    """rows `idx` of synth_dense(rows, cols, seed) without generating the whole matrix"""
    full_idx = (np.asarray(idx, dtype=np.uint64)[:, None] * np.uint64(cols) + np.arange(cols, dtype=np.uint64)[None, :])
    with np.errstate(over="ignore"):
        s = oracle._splitmix64(np.array([seed], dtype=np.uint64))[0]
        h = oracle._splitmix64(s ^ (full_idx * np.uint64(0xD1342543DE82EF95) + np.uint64(1)))
    u = (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    mag = u * (0.5 - 0.0001) + 0.0001
    return np.where((h & np.uint64(1)) != 0, -mag, mag)


def test_amazon14k_full_size_properties(nb, oracle):
    batch, feat, hid, lab, blk = 10000, 597540, 1000, 14588, 1000
    DS = nb.DeviceSet
    w1 = DS.create("ff", "w1", hid, feat, blk, blk); w1.fill_synth(2)
    b1 = DS.create("ff", "b1", hid, 1, blk, 1); b1.fill_synth(3)
    wo = DS.create("ff", "wo", lab, hid, blk, blk); wo.fill_synth(4)
    bo = DS.create("ff", "bo", lab, 1, blk, 1); bo.fill_synth(5)
    x = DS.create("ff", "inputs", batch, feat, blk, blk); x.fill_synth(1)
    out, labels, stats = nb.ff_inference_unit(w1, wo, x, b1, bo, nb.PREC_F16)
    assert out.info() == (batch, lab, blk, blk)
    # idempotence: a second run (inputs re-converted) gives the same labels
    x.invalidate_planes()
    _, labels2, _ = nb.ff_inference_unit(w1, wo, x, b1, bo, nb.PREC_F16)
    assert np.array_equal(labels, labels2)
    # a sample of rows in fp64 on the host, from the same generator (the device generator is bit-identical to the oracle's)
    sample = np.array([0, 1, 999, 1000, 4242, 7777, 9998, 9999])
    W1 = oracle.synth_dense(hid, feat, 2)                    # 4.8 GB on the host
    Xs = _rows(oracle, batch, feat, 1, sample)
    assert np.array_equal(Xs[0, :7], oracle.synth_dense(1, 7, 1)[0])
    y1 = np.maximum(Xs @ W1.T + oracle.synth_dense(hid, 1, 3).reshape(1, -1), 0.0)
    del W1
    z = y1 @ oracle.synth_dense(lab, hid, 4).T + oracle.synth_dense(lab, 1, 5).reshape(1, -1)
    assert np.array_equal(labels[sample], np.argmax(z, axis=1))
    zs = nb.DeviceSet.lookup("ff", "__logits")
    got = np.stack([np.ascontiguousarray(zs.download_block(int(r) // blk, 0, blk, lab)[int(r) % blk]) for r in sample[:3]])
    assert np.abs(got - z[:3]).max() <= 2.5e-3 * np.abs(z).max()
    # every one of the 10,000 labels against the device's own fp64 (DMMA) pipeline, which is checked against the
    # oracle at oracle-sized shapes; plus the bf16x3 path as a third, independent arithmetic
    z_f16 = zs.download_block(3, 0, blk, lab)
    _, labels_x3, _ = nb.ff_inference_unit(w1, wo, x, b1, bo, nb.PREC_BF16X3)
    assert np.array_equal(labels, labels_x3)
    _, labels_f64, _ = nb.ff_inference_unit(w1, wo, x, b1, bo, nb.PREC_FP64)
    assert np.array_equal(labels, labels_f64), f"{(labels != labels_f64).sum()} of {batch} labels differ from the fp64 pipeline"
    z_f64 = nb.DeviceSet.lookup("ff", "__logits").download_block(3, 0, blk, lab)
    assert np.abs(z_f16 - z_f64).max() <= 2.5e-3 * np.abs(z_f64).max()
    z_f64_b4 = nb.DeviceSet.lookup("ff", "__logits").download_block(4, 0, blk, lab)
    assert np.abs(z_f64_b4[242] - z[4]).max() <= 1e-9 * np.abs(z).max()   # row 4242 = sample[4]: fp64 device vs fp64 host
    out, _, _ = nb.ff_inference_unit(w1, wo, x, b1, bo, nb.PREC_F16)
    # probabilities: rows sum to one, argmax agrees with the labels, padding of the last label block is zero
    pblk = out.download_block(9, 14)                          # rows 9000..9999, labels 14000..14999 (588 real)
    assert (pblk[:, 588:] == 0).all() and (pblk >= 0).all()
    rowsum = sum(out.download_block(9, j).sum(axis=1) for j in range(15))
    assert np.abs(rowsum - 1).max() <= 1e-9
    assert stats["relevel_rows"] < batch // 10 and stats["refined_rows"] <= stats["relevel_rows"]
    for s in (w1, x):
        s.destroy()


def test_word2vec_reference_shape(nb, oracle):
    """Word2Vec.cc:106-109: vocab 1,000,000, dim 500, batch 100 (the reference's own size), 1,000 x 1,000 blocks."""
    dim, vocab, batch, blk = 500, 1000000, 100, 1000
    W = nb.DeviceSet.create("word2vec", "weights0", dim, vocab, blk, blk); W.fill_synth(21)
    X = nb.DeviceSet.create("word2vec", "inputs", batch, vocab, blk, blk); X.fill_synth(22)
    idx = np.array([0, 17, 99])
    xs = _rows(oracle, batch, vocab, 22, idx)
    wsub = _rows(oracle, dim, vocab, 21, np.arange(0, dim, 50))
    ref = wsub @ xs.T                                       # (10 sampled dims) x (3 sampled batch rows)
    for prec, tol in (("f16", 2.5e-3), ("bf16x3", 1e-4)):
        out = nb.gemm_blocks(nb.OP_NT, W, X, nb.EPI_NONE, None, nb.PRECISIONS[prec], "word2vec", "outputs")
        assert out.info() == (dim, batch, blk, blk)
        got = out.download_dense()[np.arange(0, dim, 50)][:, idx]
        assert np.abs(got - ref).max() <= tol * np.abs(ref).max() * 4
    W.destroy(), X.destroy()
