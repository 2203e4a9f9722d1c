"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on identical seeded
blocks.  Bars: bit-exact for generator / layout / index work and for predicted labels; logits within
the stated tolerance relative to the logit scale max|z|:
    FP64   1e-12   (summation order only)
    BF16X3 1e-4    (north_star: <= 1e-4 at fp32 accumulation)
    BF16   2e-2    (the looser stated bound for the single-pass bf16 variant)
    F16    2.5e-3  (single-pass fp16 with power-of-two scaling: 3 more mantissa bits than bf16)
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
TOL = {"fp64": 1e-12, "bf16x3": 1e-4, "bf16": 2e-2, "f16": 2.5e-3}


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def test_device_against_reference_goldens(nb):
    """The device (fp64 path, through the C ABI) against vectors recorded from the REFERENCE'S OWN UDF code
    (tests/golden/ref_pinned.npz, see tests/golden/make_golden_ref.py) -- no oracle in between."""
    g = np.load(os.path.join(G, "ref_pinned.npz"))
    fp64 = nb.PREC_FP64
    close = lambda a, b: np.abs(a - b).max() <= 1e-12 * max(np.abs(b).max(), 1.0)

    def dev(name, m, br=10, bc=10):
        m = np.ascontiguousarray(m, dtype=np.float64)
        s = nb.DeviceSet.create("gold", name, m.shape[0], m.shape[1], br, bc)
        s.upload_dense(m)
        return s

    w1, x, b1, wo = dev("w1", g["ff_w1"]), dev("x", g["ff_x"]), dev("b1", g["ff_b1"], 10, 1), dev("wo", g["ff_wo"])
    assert close(nb.gemm_blocks(nb.OP_NT, w1, x, nb.EPI_NONE, None, fp64, "gold", "y1pre").download_dense(), g["ff_y1pre"])
    y1 = nb.gemm_blocks(nb.OP_NT, w1, x, nb.EPI_BIAS_RELU, b1, fp64, "gold", "y1")
    assert close(y1.download_dense(), g["ff_y1"])
    assert close(nb.gemm_blocks(nb.OP_NN, wo, y1, nb.EPI_NONE, None, fp64, "gold", "zpre").download_dense(), g["ff_zpre"])
    wx = nb.gemm_blocks(nb.OP_NN, dev("lw", g["lstm_w"]), dev("lx", g["lstm_xt"]), nb.EPI_NONE, None, fp64, "gold", "wx")
    uh = nb.gemm_blocks(nb.OP_NN, dev("lu", g["lstm_u"]), dev("lh", g["lstm_h"]), nb.EPI_NONE, None, fp64, "gold", "uh")
    lb = dev("lb", g["lstm_b"])
    assert close(nb.sum_blocks(wx, uh, lb, nb.ACT_SIGMOID, "gold", "gs").download_dense(), g["lstm_gate_sigmoid"])
    assert close(nb.sum_blocks(wx, uh, lb, nb.ACT_TANH, "gold", "gt").download_dense(), g["lstm_gate_tanh"])
    assert close(nb.sum_blocks(wx, uh, None, nb.ACT_NONE, "gold", "g2").download_dense(), g["lstm_two_sum"])
    th = nb.sum_blocks(dev("lc", g["lstm_c"]), None, None, nb.ACT_TANH, "gold", "th")
    assert close(nb.gemm_blocks(nb.OP_NN, dev("lo", g["lstm_w"][:10, :20]), th, nb.EPI_NONE, None, fp64, "gold", "hid").download_dense(), g["lstm_hidden"])
    la_a, la_b = dev("la", g["la_a"]), dev("lbb", g["la_b"])
    assert close(nb.gemm_blocks(nb.OP_TN, la_a, la_a, nb.EPI_NONE, None, fp64, "gold", "gram").download_dense(), g["la_gram"])
    assert close(nb.gemm_blocks(nb.OP_TN, la_a, la_b, nb.EPI_NONE, None, fp64, "gold", "tn").download_dense(), g["la_tn"])
    assert close(nb.gemm_blocks(nb.OP_NN, dev("lat", g["la_a"].T.copy()), la_b, nb.EPI_NONE, None, fp64, "gold", "nn").download_dense(), g["la_nn"])
    assert np.array_equal(nb.conv2d_image_flat(g["conv_imgs"], 3, 1, 0, 7, 5).download_dense(), g["conv_image_flat"])
    assert np.array_equal(nb.conv2d_image_flat(g["conv_imgs2"], 3, 2, 1, 16, 4).download_dense(), g["conv_image_flat2"])
    kf = nb.conv2d_kernel_flat(g["conv_kern"], np.zeros(5), 7, 5)
    assert np.array_equal(kf.download_dense(), g["conv_kernel_flat_nobias"])
    res = nb.conv2d_fused(g["conv_imgs"], g["conv_kern"], np.zeros(5), 0, False, fp64, 7, 5)
    assert close(res.download_dense(), g["conv_result_nobias"])
    one = dev("one", g["conv_one_result"], 7, 2)
    assert np.array_equal(nb.conv2d_result_to_images(one, 1, 4, 6, 5), g["conv_one_images"])
    # FullyConnectedNetwork / SemanticClassifier: the output set holds the labels x batch blocks side by side
    fc = nb.udf_dense2(dev("fcx", g["fc_x"], 6, 10), 10, 6, g["fc_w0"], g["fc_b0"], g["fc_w1"], g["fc_b1"], "gold", "fc").download_dense()
    assert np.array_equal(fc.reshape(3, 3, 2, 6).transpose(0, 2, 1, 3).reshape(6, 3, 6), g["fc_out"])
    sc = nb.udf_dense2(dev("scx", g["sc_x"], 12, 5), 12, 5, np.ones((4, 12)), np.ones(4), np.ones((2, 4)), np.ones(2), "gold", "sc").download_dense()
    assert np.array_equal(sc.reshape(1, 2, 3, 5).transpose(0, 2, 1, 3).reshape(3, 2, 5), g["sc_out"])


def test_generator_bit_exact(nb, oracle):
    s = nb.DeviceSet.create("t", "gen", 37, 53, 10, 10)
    s.fill_synth(42)
    assert np.array_equal(s.download_dense(), oracle.synth_dense(37, 53, 42))
    s.fill_synth(42, row_offset=5)  # a row shard reproduces its slice of the larger matrix
    assert np.array_equal(s.download_dense(), oracle.synth_dense(42, 53, 42)[5:])
    s.destroy()


def test_block_upload_download_roundtrip(nb, oracle):
    """sendData / getSetIterator: full-size zero-padded edge blocks in, zero-padded blocks out."""
    d = oracle.synth_dense(23, 17, 3)
    ref = oracle.load_matrix(d, 10, 6)
    s = nb.DeviceSet.create("t", "rt", 23, 17, 10, 6)
    for br, bc, _, _, arr in ref.blocks():
        s.upload_block(br, bc, np.ascontiguousarray(arr))
    nb.synchronize()
    assert np.array_equal(s.download_dense(), d)
    for br, bc, _, _, arr in ref.blocks():
        assert np.array_equal(s.download_block(br, bc), arr)
    # trimmed (dont_pad) blocks and an all-padding block
    t = nb.DeviceSet.create("t", "rt2", 23, 17, 10, 6)
    for br, bc, _, _, arr in oracle.load_matrix(d, 10, 6, True, True).blocks():
        t.upload_block(br, bc, np.ascontiguousarray(arr))
    t.upload_block(5, 5, np.ones((10, 6)))
    nb.synchronize()
    assert np.array_equal(t.download_dense(), d)
    s.destroy(), t.destroy()


@pytest.mark.parametrize("prec", ["fp64", "bf16x3", "bf16", "f16"])
@pytest.mark.parametrize("shape", [(230, 310, 150, 100), (64, 128, 64, 64), (5, 7, 3, 4), (300, 2500, 129, 100),
                                   (12, 20002, 300, 1000), (16, 8222, 129, 100), (9, 9001, 40, 1000)])  # skinny: the streaming fp64 kernel
def test_transpose_mult_and_input_layer_join(nb, oracle, prec, shape):
    I, K, J, blk = shape
    a, b = oracle.synth_dense(I, K, 11), oracle.synth_dense(J, K, 12)
    A = nb.DeviceSet.create("t", "a", I, K, blk, blk); A.upload_dense(a)
    B = nb.DeviceSet.create("t", "b", J, K, blk, blk); B.upload_dense(b)
    C = nb.gemm_blocks(nb.OP_NT, A, B, nb.EPI_NONE, None, nb.PRECISIONS[prec], "t", "c")
    ref = oracle.transpose_mult_agg(oracle.load_matrix(a, blk, blk), oracle.load_matrix(b, blk, blk), 2)
    assert C.info() == (I, J, blk, blk) == ref.totals() + (blk, blk)   # FFTransposeMult.h:68-71
    assert rel(C.download_dense(), ref.dense()) <= TOL[prec]
    # blocks come back zero padded like the reference's
    for br, bc, _, _, arr in ref.blocks():
        got = C.download_block(br, bc)
        assert got.shape == arr.shape and np.abs(got - arr).max() <= TOL[prec] * np.abs(ref.dense()).max()
    # FFInputLayerJoin: in1(I x K) . in2(K x J)
    bt = np.ascontiguousarray(b.T)
    Bt = nb.DeviceSet.create("t", "bt", K, J, blk, blk); Bt.upload_dense(bt)
    C2 = nb.gemm_blocks(nb.OP_NN, A, Bt, nb.EPI_NONE, None, nb.PRECISIONS[prec], "t", "c2")
    ref2 = oracle.input_layer_join_agg(oracle.load_matrix(a, blk, blk), oracle.load_matrix(bt, blk, blk), 2)
    assert C2.info()[:2] == ref2.totals()
    assert rel(C2.download_dense(), ref2.dense()) <= TOL[prec]


@pytest.mark.parametrize("prec", ["fp64", "bf16x3", "bf16", "f16"])
def test_fused_bias_epilogues(nb, oracle, prec):
    I, K, J, blk = 130, 270, 90, 50
    a, b, bias = oracle.synth_dense(I, K, 1), oracle.synth_dense(J, K, 2), oracle.synth_dense(I, 1, 3) * 20
    A = nb.DeviceSet.create("t", "a", I, K, blk, blk); A.upload_dense(a)
    B = nb.DeviceSet.create("t", "b", J, K, blk, blk); B.upload_dense(b)
    Bi = nb.DeviceSet.create("t", "bias", I, 1, blk, 1); Bi.upload_dense(bias)
    oa, ob, obias = oracle.load_matrix(a, blk, blk), oracle.load_matrix(b, blk, blk), oracle.load_matrix(bias, blk, 1)
    pre = oracle.transpose_mult_agg(oa, ob, 1)
    scale = np.abs(pre.dense()).max()
    relu = nb.gemm_blocks(nb.OP_NT, A, B, nb.EPI_BIAS_RELU, Bi, nb.PRECISIONS[prec], "t", "relu")
    ref = oracle.relu_bias_sum(pre, obias).dense()
    got = relu.download_dense()
    assert np.abs(got - ref).max() <= TOL[prec] * scale and (got >= 0).all() and (got == 0).any()
    tr = nb.gemm_blocks(nb.OP_NT, A, B, nb.EPI_BIAS_TRANSPOSE, Bi, nb.PRECISIONS[prec], "t", "tr")
    reft = oracle.transpose_bias_sum(pre, obias, apply_exp=False)
    assert tr.info() == (J, I, blk, blk) == reft.totals() + (blk, blk)  # FFTransposeBiasSum.h:64-67
    assert np.abs(tr.download_dense() - reft.dense()).max() <= TOL[prec] * scale


def test_bias_shape_mismatch_is_reported(nb, oracle):
    A = nb.DeviceSet.create("t", "a", 8, 8, 4, 4); A.fill_synth(1)
    B = nb.DeviceSet.create("t", "b", 8, 8, 4, 4); B.fill_synth(2)
    bad = nb.DeviceSet.create("t", "bad", 8, 2, 4, 2); bad.fill_synth(3)
    with pytest.raises(nb.Nb200Error, match="mismatch"):
        nb.gemm_blocks(nb.OP_NT, A, B, nb.EPI_BIAS_RELU, bad, nb.PREC_BF16X3, "t", "x")
    B2 = nb.DeviceSet.create("t", "b2", 8, 8, 4, 2); B2.fill_synth(2)  # block K differs (FFTransposeMult.h:61-65)
    with pytest.raises(nb.Nb200Error, match="mismatch"):
        nb.gemm_blocks(nb.OP_NT, A, B2, nb.EPI_NONE, None, nb.PREC_BF16X3, "t", "x")


def test_softmax_rows_vs_row_aggregate_output_layer(nb, oracle):
    z = oracle.synth_dense(57, 301, 8) * 40      # |z| up to 20: exp stays finite in the oracle
    Z = nb.DeviceSet.create("t", "z", 57, 301, 25, 100); Z.upload_dense(z)
    P, labels = nb.softmax_rows(Z, "t", "p")
    yo = oracle.load_matrix(np.exp(z), 25, 100)
    ref = oracle.output_layer(yo, oracle.row_aggregate(yo)).dense()
    assert np.abs(P.download_dense() - ref).max() <= 1e-14
    assert np.array_equal(labels, np.argmax(z, axis=1))
    # first maximum wins on exact ties
    z[3, 7] = z[3, 200] = 99.0
    Z.upload_dense(z)
    _, labels = nb.softmax_rows(Z, "t", "p")
    assert labels[3] == 7


def _ff_sets(nb, oracle, batch, feat, hid, lab, blk, wo_scale=1.0, dense=None):
    if dense is None:
        dense = dict(x=oracle.synth_dense(batch, feat, 1), w1=oracle.synth_dense(hid, feat, 2),
                     b1=oracle.synth_dense(hid, 1, 3), wo=oracle.synth_dense(lab, hid, 4) * wo_scale,
                     bo=oracle.synth_dense(lab, 1, 5))
    dims = dict(x=(batch, feat, blk, blk), w1=(hid, feat, blk, blk), b1=(hid, 1, blk, 1), wo=(lab, hid, blk, blk),
                bo=(lab, 1, blk, 1))
    names = dict(x="inputs", w1="w1", b1="b1", wo="wo", bo="bo")
    dev, ora = {}, {}
    for k, (r, c, br, bc) in dims.items():
        dev[k] = nb.DeviceSet.create("ff", names[k], r, c, br, bc)
        dev[k].upload_dense(dense[k])
        ora[k] = oracle.load_matrix(dense[k], br, bc)
    return dev, ora, dense


@pytest.mark.parametrize("prec", ["fp64", "bf16x3", "bf16", "f16"])
def test_ff_inference_unit_config1(nb, oracle, prec):
    """BASELINE config 1: batch 1,000 x 1,024 features, hidden 1,024, 10 labels, 100 x 100 blocks."""
    dev, ora, _ = _ff_sets(nb, oracle, 1000, 1024, 1024, 10, 100)
    out, labels, stats = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PRECISIONS[prec])
    ref_out, ref_logits, ref_labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], ora["x"], ora["b1"], ora["bo"], threads=4)
    assert np.array_equal(labels, ref_labels), f"{(labels != ref_labels).sum()} labels differ"
    assert out.info() == (1000, 10, 100, 100) == ref_out.totals() + (100, 100)
    z = nb.DeviceSet.lookup("ff", "__logits").download_dense()
    zr = ref_logits.dense()
    assert rel(z, zr) <= TOL[prec], rel(z, zr)
    p, pr = out.download_dense(), ref_out.dense()
    assert np.isfinite(pr).all()
    assert np.abs(p.sum(axis=1) - 1).max() <= 1e-12
    assert np.abs(p - pr).max() <= 4 * TOL[prec] * np.abs(zr).max() + 1e-13
    # the `yo` set of the reference (exp of the transposed, biased logits) is a view on the device
    yo = nb.DeviceSet.lookup("ff", "yo").download_dense()
    assert np.allclose(yo, np.exp(z), rtol=1e-12)
    assert stats["kernel_launches"] > 0


def test_ff_golden_fixtures(nb, oracle):
    g = np.load(os.path.join(G, "kat_2x3.npz"))
    dense = {k: g[k] for k in ("x", "w1", "b1", "wo", "bo")}
    for prec in ("fp64", "bf16x3", "bf16", "f16"):
        dev, _, _ = _ff_sets(nb, oracle, 2, 3, 2, 2, int(g["block"]), dense=dense)
        out, labels, _ = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PRECISIONS[prec])
        assert np.array_equal(labels, g["labels"])
        # every value of this KAT is exactly representable in bf16, so all precisions agree to fp64 rounding
        assert np.abs(out.download_dense() - g["probs"]).max() <= 1e-6 if prec != "fp64" else 1e-15
    g = np.load(os.path.join(G, "ff_small.npz"))
    batch, feat, hid, lab, blk = (int(g[k]) for k in ("batch", "feat", "hid", "lab", "block"))
    for prec in ("fp64", "bf16x3", "bf16", "f16"):
        dev, _, _ = _ff_sets(nb, oracle, batch, feat, hid, lab, blk, wo_scale=float(g["wo_scale"]))
        out, labels, _ = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PRECISIONS[prec])
        assert np.array_equal(labels, g["labels"])
        z = nb.DeviceSet.lookup("ff", "__logits").download_dense()
        assert rel(z, g["logits"]) <= TOL[prec]


def test_ff_ragged_and_tiny_and_empty(nb, oracle):
    for (batch, feat, hid, lab, blk) in [(1, 1, 1, 1, 1), (3, 5, 2, 4, 2), (129, 65, 257, 33, 64), (130, 1000, 70, 300, 100)]:
        dev, ora, _ = _ff_sets(nb, oracle, batch, feat, hid, lab, blk)
        ref_out, ref_logits, ref_labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], ora["x"], ora["b1"], ora["bo"], threads=2)
        for prec in ("fp64", "bf16x3", "bf16", "f16"):
            out, labels, _ = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PRECISIONS[prec])
            assert np.array_equal(labels, ref_labels), (batch, feat, hid, lab, prec)
            z = nb.DeviceSet.lookup("ff", "__logits").download_dense()
            assert rel(z, ref_logits.dense()) <= TOL[prec]
    # empty batch: nothing to do, no error
    w1 = nb.DeviceSet.lookup("ff", "w1"); wo = nb.DeviceSet.lookup("ff", "wo")
    b1 = nb.DeviceSet.lookup("ff", "b1"); bo = nb.DeviceSet.lookup("ff", "bo")
    empty = nb.DeviceSet.create("ff", "inputs", 0, 1000, 100, 100); empty.fill_synth(1)
    out, labels, _ = nb.ff_inference_unit(w1, wo, empty, b1, bo, nb.PREC_BF16X3)
    assert out.info()[0] == 0 and labels.shape == (0,)


def test_near_tie_rows_are_re_evaluated(nb, oracle):
    """Two identical label rows in wo make every row an exact top-2 tie candidate for those labels: the
    low-precision path must flag and re-evaluate, and still return the first maximum like the oracle."""
    batch, feat, hid, lab, blk = 200, 640, 128, 12, 64
    dense = dict(x=oracle.synth_dense(batch, feat, 1), w1=oracle.synth_dense(hid, feat, 2),
                 b1=oracle.synth_dense(hid, 1, 3), wo=oracle.synth_dense(lab, hid, 4), bo=oracle.synth_dense(lab, 1, 5))
    dense["wo"][7] = dense["wo"][2] * (1 + 1e-9)   # nearly identical logits for labels 2 and 7
    dense["bo"][7] = dense["bo"][2]
    dev, ora, _ = _ff_sets(nb, oracle, batch, feat, hid, lab, blk, dense=dense)
    _, _, ref_labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], ora["x"], ora["b1"], ora["bo"], threads=2)
    for prec in ("bf16x3", "bf16", "f16"):
        _, labels, stats = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PRECISIONS[prec])
        assert np.array_equal(labels, ref_labels)
        assert stats["refined_rows"] >= int(np.isin(ref_labels, (2, 7)).sum())


def test_f16_scaling_handles_large_and_tiny_magnitudes(nb, oracle):
    """fp16 has 5 exponent bits; the per-matrix power-of-two scales must keep products of very large or very small
    inputs inside range, and an input outlier that the sampled scale misses must fall back (not return inf)."""
    I, K, J, blk = 150, 900, 140, 100
    a, b = oracle.synth_dense(I, K, 31), oracle.synth_dense(J, K, 32)
    for sa, sb in ((1e6, 3e5), (1e-6, 2e-7), (1e5, 1e-5)):
        A = nb.DeviceSet.create("t", "a", I, K, blk, blk); A.upload_dense(a * sa)
        B = nb.DeviceSet.create("t", "b", J, K, blk, blk); B.upload_dense(b * sb)
        C = nb.gemm_blocks(nb.OP_NT, A, B, nb.EPI_NONE, None, nb.PREC_F16, "t", "c").download_dense()
        assert np.isfinite(C).all() and rel(C, (a * sa) @ (b * sb).T) <= TOL["f16"]
    # inference with one huge input element in a row that the 1-in-64 row sample does not see
    batch, feat, hid, lab = 200, 300, 64, 7
    dense = dict(x=oracle.synth_dense(batch, feat, 1), w1=oracle.synth_dense(hid, feat, 2), b1=oracle.synth_dense(hid, 1, 3),
                 wo=oracle.synth_dense(lab, hid, 4), bo=oracle.synth_dense(lab, 1, 5))
    dense["x"][37, 11] = 4.0e3          # 8000x the typical magnitude: beyond the 32x headroom of the sampled scale
    dev, ora, _ = _ff_sets(nb, oracle, batch, feat, hid, lab, 50, dense=dense)
    _, ref_logits, ref_labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], ora["x"], ora["b1"], ora["bo"], threads=2)
    _, labels, _ = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PREC_F16)
    z = nb.DeviceSet.lookup("ff", "__logits").download_dense()
    assert np.isfinite(z).all() and np.array_equal(labels, ref_labels)
    assert rel(z, ref_logits.dense()) <= TOL["bf16"]


def test_conv2d_memory_fusion(nb, oracle):
    g = np.load(os.path.join(G, "conv_small.npz"))
    IF = nb.conv2d_image_flat(g["images"], 3, 1, 0, 4, 5)
    assert np.array_equal(IF.download_dense(), g["image_flat"])           # pure data movement: bit exact
    KF = nb.conv2d_kernel_flat(g["kernels"], g["bias"], 4, 5)
    assert np.array_equal(KF.download_dense(), g["kernel_flat"])
    for prec in ("fp64", "bf16x3", "bf16", "f16"):
        R = nb.gemm_blocks(nb.OP_NT, IF, KF, nb.EPI_NONE, None, nb.PRECISIONS[prec], "conv2d", "result")
        assert rel(R.download_dense(), g["result"]) <= TOL[prec]
        F = nb.conv2d_fused(g["images"], g["kernels"], g["bias"], 0, True, nb.PRECISIONS[prec], 4, 5)
        assert np.abs(F.download_dense() - np.maximum(g["result"], 0)).max() <= TOL[prec] * np.abs(g["result"]).max()
    # larger, padded / strided im2col against the oracle's blocked image_flat
    imgs = oracle.synth_dense(3 * 8, 20 * 19, 5).reshape(3, 8, 20, 19)
    for k, s, p in ((3, 1, 0), (3, 2, 1), (5, 1, 2)):
        ref = oracle.image_flat(imgs, k, s, p, 32, 32)
        got = nb.conv2d_image_flat(imgs, k, s, p, 32, 32)
        assert got.info() == ref.totals() + (32, 32)
        assert np.array_equal(got.download_dense(), ref.dense())
        for br, bc, _, _, arr in list(ref.blocks())[:6]:
            assert np.array_equal(got.download_block(br, bc), arr)


def test_word2vec_block_join(nb, oracle):
    """Word2Vec.cc:53-71 at reduced size: weights (dim x vocab) x inputs (batch x vocab) by FFTransposeMult."""
    dim, vocab, batch, blk = 50, 20000, 300, 1000
    w, x = oracle.synth_dense(dim, vocab, 21), oracle.synth_dense(batch, vocab, 22)
    W = nb.DeviceSet.create("word2vec", "weights0", dim, vocab, blk, blk); W.upload_dense(w)
    X = nb.DeviceSet.create("word2vec", "inputs", batch, vocab, blk, blk); X.upload_dense(x)
    ref = oracle.transpose_mult_agg(oracle.load_matrix(w, blk, blk), oracle.load_matrix(x, blk, blk), 4).dense()
    for prec in ("fp64", "bf16x3", "bf16", "f16"):
        out = nb.gemm_blocks(nb.OP_NT, W, X, nb.EPI_NONE, None, nb.PRECISIONS[prec], "word2vec", "outputs")
        assert out.info()[:2] == (dim, batch)
        assert rel(out.download_dense(), ref) <= TOL[prec]


def _run(cmd, env_extra, timeout=300):
    import subprocess

    env = dict(os.environ, **env_extra)
    return subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)


@pytest.mark.parametrize("prec", ["bf16", "f16", "bf16x3", "fp64"])
def test_cpp_driver_through_netsdb_api(oracle, prec, tmp_path):
    """The C++ host shim end to end: FFTest builds the reference's Computation graphs (FFTransposeMult ->
    FFAggMatrix -> FFReluBiasSum -> FFInputLayerJoin -> FFAggMatrix -> FFTransposeBiasSum; FFRowAggregate ->
    FFOutputLayer) through PDBClient::executeComputations and reads `output` with getSetIterator."""
    from netsdb_b200 import build

    exe = build.build_host()["FFTest"]
    blk, batch, feat, hid, lab = 100, 250, 640, 300, 12
    dump = tmp_path / "dump.txt"
    r = _run([exe, str(blk), str(blk), str(batch), str(feat), str(hid), str(lab), str(dump)], {"NB200_PRECISION": prec})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    got = np.loadtxt(dump)
    # the shim's loader uses seeds inputs 1, w1 2, b1 3, wo 4, bo 5 and the reference driver's block shapes
    X = oracle.load_matrix(oracle.synth_dense(batch, feat, 1), blk, blk)
    W1 = oracle.load_matrix(oracle.synth_dense(hid, feat, 2), blk, blk)
    B1 = oracle.load_matrix(oracle.synth_dense(hid, 1, 3), blk, 1)
    WO = oracle.load_matrix(oracle.synth_dense(lab, hid, 4), blk, blk)
    BO = oracle.load_matrix(oracle.synth_dense(lab, 1, 5), blk, 1)
    ref_out, _, ref_labels = oracle.ff_inference_unit(W1, WO, X, B1, BO, threads=2)
    assert np.array_equal(got[:, 0].astype(np.int32), ref_labels)          # fused stage's labels
    assert np.array_equal(got[:, 1].astype(np.int32), ref_labels)          # argmax of the `output` set read back
    assert np.abs(got[:, 2] - ref_out.dense()[:, 0]).max() <= 4 * TOL[prec] * 200 + 1e-13


def test_unmodified_reference_driver_runs(tmp_path):
    """netsDB's own src/tests/source/FFTest.cc, compiled unmodified against the shim (only where the reference
    tree was present at build time), runs against the device.  It asserts nothing itself;
    exit status and its block count line are what can be checked."""
    from netsdb_b200 import build

    exe = os.path.join(build.HOST_BIN, "FFTest_netsdb")
    if not os.path.exists(exe):
        pytest.skip("reference driver was not built (no reference tree at build time)")
    # blockX blockY batch features neurons labels reload; the driver passes dropout 0.5 (time-seeded in netsDB)
    r = _run([exe, "50", "50", "120", "200", "100", "2", "Y"], {"NB200_DROPOUT_AS_ZERO": "1"}, timeout=120)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert "count=3" in r.stdout  # 120 rows / 50 = 3 row blocks x 1 column block of the output set


def test_unmodified_reference_driver_with_sizes():
    """netsDB's src/tests/source/FFTestWithoutDeduplication.cc (sizes, database name and reload flag from the command
    line; the driver bench.py's configuration can be run with) unmodified through the shim."""
    from netsdb_b200 import build

    exe = os.path.join(build.HOST_BIN, "FFTestWithoutDeduplication_netsdb")
    if not os.path.exists(exe):
        pytest.skip("reference driver was not built (no reference tree at build time)")
    r = _run([exe, "50", "50", "130", "220", "90", "7", "ffnd", "Y"], {"NB200_DROPOUT_AS_ZERO": "1", "NB200_PRECISION": "f16"}, timeout=120)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert "*****FFTest End-to-End Time Duration" in r.stdout


def test_upload_commit_overlaps_next_batch(nb, oracle):
    """Double-buffered inputs: the blocks of batch B are queued (and committed) while batch A is evaluated, the results
    of A are downloaded while B's transfer may still be running, then B is evaluated.  Labels and probabilities of both
    batches equal the oracle's, repeatedly, and a late upload withdraws the commit."""
    batch, feat, hid, lab, blk = 1200, 8200, 300, 20, 1000
    dev, ora, dense = _ff_sets(nb, oracle, batch, feat, hid, lab, blk)
    xb = oracle.synth_dense(batch, feat, 77)
    ref = {}
    for name, xd in (("a", dense["x"]), ("b", xb)):
        out, _, labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], oracle.load_matrix(xd, blk, blk), ora["b1"], ora["bo"], threads=4)
        ref[name] = (labels, out.dense())
    A, B = dev["x"], nb.DeviceSet.create("ff", "inputs_b", batch, feat, blk, blk)
    host = {"a": oracle.load_matrix(dense["x"], blk, blk), "b": oracle.load_matrix(xb, blk, blk)}

    def send(dst, name):
        for br, bc, _, _, arr in host[name].blocks():
            dst.upload_block(br, bc, np.ascontiguousarray(arr))
        dst.upload_commit()

    send(A, "a")
    for rep in range(3):
        send(B, "b")                                                    # in flight while A is evaluated
        out, labels, _ = nb.ff_inference_unit(dev["w1"], dev["wo"], A, dev["b1"], dev["bo"], nb.PREC_F16)
        assert np.array_equal(labels, ref["a"][0]), rep
        assert np.abs(out.download_dense() - ref["a"][1]).max() < 5e-2  # fp16 logits; the labels above are the exact check
        send(A, "a")                                                    # in flight while B is evaluated
        out, labels, _ = nb.ff_inference_unit(dev["w1"], dev["wo"], B, dev["b1"], dev["bo"], nb.PREC_F16)
        assert np.array_equal(labels, ref["b"][0]), rep
        assert np.abs(out.download_dense() - ref["b"][1]).max() < 5e-2
    # an upload after the commit withdraws it: the reader falls back to waiting for everything and sees the new data
    B.upload_commit()
    for br, bc, _, _, arr in host["a"].blocks():
        B.upload_block(br, bc, np.ascontiguousarray(arr))
    _, labels, _ = nb.ff_inference_unit(dev["w1"], dev["wo"], B, dev["b1"], dev["bo"], nb.PREC_F16)
    assert np.array_equal(labels, ref["a"][0])
    B.destroy()


@pytest.mark.parametrize("trim", [False, True])
def test_fp16_side_format_inputs(nb, oracle, trim):
    """the inputs kept as an fp16 operand plane next to their fp64 blocks.  Exported once from the resident
    fp64 set, imported into a set that holds NO fp64 data on the device, evaluated with the label-exact cascade whose
    undecided rows are read straight out of the pinned host blocks (zero-padded or trimmed edge blocks).  Labels equal the
    oracle's and the fp64-resident run's; logits are bit-identical to that run's first level."""
    import torch

    batch, feat, hid, lab, blk = 1300, 8300, 300, 40, 1000
    dev, ora, dense = _ff_sets(nb, oracle, batch, feat, hid, lab, blk, wo_scale=0.02)  # small wo: many near ties
    _, _, ref_labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], ora["x"], ora["b1"], ora["bo"], threads=4)
    nb.set_option("near_tie_tau_x10", 80)  # a wide band: more than 16 undecided rows, so that BOTH levels fetch rows from the host
    plane, e, ss = dev["x"].export_f16()
    _, labels_res, st_res = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PREC_F16)
    z_res = nb.DeviceSet.lookup("ff", "__logits").download_dense()
    assert np.array_equal(labels_res, ref_labels)
    # the fp64 blocks in pinned host memory, as a netsDB page would hold them
    n_rb, n_cb = -(-batch // blk), -(-feat // blk)
    pinned, ptrs = [], np.zeros(n_rb * n_cb, dtype=np.uint64)
    for i in range(n_rb):
        for j in range(n_cb):
            rows = min(blk, batch - i * blk) if trim else blk
            cols = min(blk, feat - j * blk) if trim else blk
            t = torch.zeros(rows * cols, dtype=torch.float64).pin_memory()
            b = t.numpy().reshape(rows, cols)
            src = dense["x"][i * blk:i * blk + rows, j * blk:j * blk + cols]
            b[:src.shape[0], :src.shape[1]] = src
            pinned.append(t)
            ptrs[i * n_cb + j] = b.ctypes.data
    edge_ld = (feat - (n_cb - 1) * blk) if trim else blk
    S = nb.DeviceSet.create("ff", "inputs_side", batch, feat, blk, blk)
    pp = torch.from_numpy(plane.view(np.int16)).pin_memory().numpy().view(np.uint16)
    for rep in range(2):
        S.import_f16(pp, e, ss, ptrs, edge_ld)
        S.upload_commit()
        out, labels, st = nb.ff_inference_unit(dev["w1"], dev["wo"], S, dev["b1"], dev["bo"], nb.PREC_F16)
        assert np.array_equal(labels, ref_labels), (labels != ref_labels).sum()
        assert st["relevel_rows"] == st_res["relevel_rows"] > 16  # the cascade ran, on rows fetched from the host blocks
        assert st["refined_rows"] == st_res["refined_rows"]
        assert np.array_equal(nb.DeviceSet.lookup("ff", "__logits").download_dense(), z_res)
    with pytest.raises(nb.Nb200Error):  # no fp64 on the device: only the fp16 evaluation is possible
        nb.ff_inference_unit(dev["w1"], dev["wo"], S, dev["b1"], dev["bo"], nb.PREC_BF16X3)
    unpinned = np.zeros(n_rb * n_cb, dtype=np.uint64) + np.uint64(np.zeros(8).ctypes.data)
    with pytest.raises(nb.Nb200Error):  # host blocks the device cannot see
        S.import_f16(pp, e, ss, unpinned, edge_ld)
    S.destroy()
    nb.set_option("near_tie_tau_x10", 25)


def test_concurrent_callers_are_serialised(nb, oracle):
    """netsDB runs one pipeline thread per hash partition, all calling UDFs at once (threading).  The
    library serialises its entry points, so concurrent calls from several threads give the same results as sequential
    ones (ctypes releases the GIL during the calls)."""
    import threading

    I, K, J, blk = 300, 2500, 129, 100
    mats = [(oracle.synth_dense(I, K, 50 + t), oracle.synth_dense(J, K, 60 + t)) for t in range(4)]
    sets = []
    for t, (a, b) in enumerate(mats):
        A = nb.DeviceSet.create("thr", f"a{t}", I, K, blk, blk); A.upload_dense(a)
        B = nb.DeviceSet.create("thr", f"b{t}", J, K, blk, blk); B.upload_dense(b)
        sets.append((A, B))
    out, errs = [None] * 4, []

    def work(t):
        try:
            for rep in range(3):
                C = nb.gemm_blocks(nb.OP_NT, sets[t][0], sets[t][1], nb.EPI_NONE, None, nb.PREC_BF16X3 if t % 2 else nb.PREC_FP64,
                                   "thr", f"c{t}")
                out[t] = C.download_dense()
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    threads = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    [th.start() for th in threads]
    [th.join() for th in threads]
    assert not errs, errs
    for t, (a, b) in enumerate(mats):
        want = a @ b.T
        assert np.abs(out[t] - want).max() <= (TOL["bf16x3"] if t % 2 else TOL["fp64"]) * np.abs(want).max(), t


def test_k_sharded_variant_two_gpus():
    """K-sharded layer 1 + NCCL reduce-scatter (the fallback for weights larger than one GPU's HBM) against the oracle;
    needs two visible GPUs, so it is skipped on the single-GPU test box (run under `gpurun --gpus 2`)."""
    import sys

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
              "--master-port", "29547", os.path.join(root, "tests", "tools", "ksharded_check.py")], {}, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("OK  k-sharded") == 3


def test_unmodified_reference_word2vec_driver(tmp_path):
    """netsDB's own src/word2vec/source/Word2Vec.cc compiled unmodified against the shim: loads `inputs` and
    `weights0` from text files (ff::load_matrix_data), runs FFTransposeMult + FFAggMatrix through
    executeComputations(errMsg, jobName, materializeHash, sink) and reports the shape of `outputs`."""
    from netsdb_b200 import build

    exe = os.path.join(build.HOST_BIN, "Word2Vec_netsdb")
    if not os.path.exists(exe):
        pytest.skip("reference driver was not built (no reference tree at build time)")
    rng = np.random.default_rng(3)
    for name, shape in (("inputs", (30, 400)), ("weights0", (20, 400))):
        with open(tmp_path / f"{name}.txt", "w") as f:
            f.write(f"{shape[0]}, {shape[1]}\n")
            np.savetxt(f, rng.uniform(-1, 1, shape), fmt="%.17g")
    r = _run([exe, "50", "50", "Y", "1", str(tmp_path)], {"NB200_PRECISION": "bf16x3"}, timeout=180)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert "[STATS] outputs : 20 X 30" in r.stdout and "count = 1" in r.stdout  # embedding_dimension x batch, one block


@pytest.mark.parametrize("shape", [(2, 3, 20, 19, 5, 3, 0), (3, 8, 17, 150, 40, 3, 1), (2, 64, 30, 30, 64, 3, 0),
                                   (1, 70, 12, 140, 130, 5, 2), (2, 3, 16, 16, 7, 7, 3)])
def test_conv2d_implicit_gemm_vs_blocked_oracle(nb, oracle, shape):
    """nb200_conv2d_fused (implicit GEMM, image_flat never built) against the reference pipeline restated by the
    oracle: image_flat x kernel_flat by FFTransposeMult + FFAggMatrix, then ReLU (config C4's "+ ReLU")."""
    n, c, h, w, kout, k, pad = shape
    imgs = oracle.synth_dense(n * c, h * w, 41).reshape(n, c, h, w)
    kern = oracle.synth_dense(kout * c, k * k, 42).reshape(kout, c, k, k)
    bias = oracle.synth_dense(kout, 1, 43).reshape(kout)
    ref = oracle.transpose_mult_agg(oracle.image_flat(imgs, k, 1, pad, 64, 64), oracle.kernel_flat(kern, bias, 64, 64), 4)
    refd = ref.dense()
    hout, wout = h + 2 * pad - k + 1, w + 2 * pad - k + 1
    assert ref.totals() == (n * hout * wout, kout)
    for prec in ("fp64", "bf16x3", "bf16"):
        for relu in (False, True):
            R = nb.conv2d_fused(imgs, kern, bias, pad, relu, nb.PRECISIONS[prec], 64, 64)
            assert R.info()[:2] == (n * hout * wout, kout)
            want = np.maximum(refd, 0) if relu else refd
            assert np.abs(R.download_dense() - want).max() <= TOL[prec] * np.abs(refd).max(), (prec, relu)


@pytest.mark.parametrize("shape", [(2, 3, 20, 19, 5, 3, 0, 5), (3, 8, 17, 150, 40, 3, 1, 40), (2, 4, 9, 9, 3, 3, 0, 6)])
def test_conv2d_result_to_images(nb, oracle, shape):
    """ConvResultToChunks + ConvChunksToImage: the result matrix re-assembled into images (pure data movement: bit
    exact against the oracle's chunk-by-chunk restatement and against the plain NCHW transpose)."""
    n, c, h, w, kout, k, pad, z = shape
    imgs = oracle.synth_dense(n * c, h * w, 41).reshape(n, c, h, w)
    kern = oracle.synth_dense(kout * c, k * k, 42).reshape(kout, c, k, k)
    bias = oracle.synth_dense(kout, 1, 43).reshape(kout)
    hout, wout = h + 2 * pad - k + 1, w + 2 * pad - k + 1
    ref = oracle.transpose_mult_agg(oracle.image_flat(imgs, k, 1, pad, 50, 64), oracle.kernel_flat(kern, bias, 64, 64), 4)
    want = oracle.conv_result_to_images(ref, n, hout, wout, z)
    R = nb.conv2d_fused(imgs, kern, bias, pad, False, nb.PREC_FP64, 50, 64)
    got = nb.conv2d_result_to_images(R, n, hout, wout, z)
    assert got.shape == (n, z, hout, wout)
    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
    # the movement itself is exact: the device result matrix, transposed on the host, equals what the call returns
    dense = R.download_dense()
    plain = np.zeros((n, z, hout * wout))
    plain[:, :kout] = dense.reshape(n, hout * wout, kout).transpose(0, 2, 1)
    assert np.array_equal(got.reshape(n, z, -1), plain)
    F = nb.conv2d_fused(imgs, kern, bias, pad, True, nb.PREC_BF16, 50, 64)  # fp32 result set
    gf = nb.conv2d_result_to_images(F, n, hout, wout, z)
    assert np.array_equal(gf.reshape(n, z, -1)[:, :kout], F.download_dense().reshape(n, hout * wout, kout).transpose(0, 2, 1))
    with pytest.raises(nb.Nb200Error):
        nb.conv2d_result_to_images(R, n, hout + 1, wout, z)
    if kout > 1:
        with pytest.raises(nb.Nb200Error):
            nb.conv2d_result_to_images(R, n, hout, wout, kout - 1)


@pytest.mark.parametrize("cfg", [(2, 3, 20, 19, 5, 3, 1, 0, 32, 32, "fp64"), (3, 4, 12, 12, 6, 3, 2, 1, 16, 10, "fp64"),
                                 (2, 8, 17, 30, 40, 3, 1, 1, 64, 64, "bf16x3")])
def test_cpp_conv2d_driver_through_netsdb_api(oracle, tmp_path, cfg):
    """The C++ host shim for conv2d_memory_fusion: Conv2dMemFuseTest builds the reference's four jobs from its
    Computation classes (KernelToChunks / ImageToChunks -> ImageChunksToBlock -> ImageBlockToMatrix [-> KernelBiasJoin],
    FFTransposeMult + FFAggMatrix, ConvResultToChunks -> ImageChunksToBlock -> ConvChunksToImage) and the shim lowers
    each job to the device; the re-assembled images are compared with the oracle's pipeline on the same data."""
    from netsdb_b200 import build

    exe = build.build_host()["Conv2dMemFuseTest"]
    n, c, h, w, kout, k, stride, pad, bx, by, prec = cfg
    r = _run([exe] + [str(v) for v in (n, c, h, w, kout, k, stride, pad, bx, by)] + [str(tmp_path)], {"NB200_PRECISION": prec}, timeout=120)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    imgs = oracle.synth_dense(n * c, h * w, 41).reshape(n, c, h, w)
    kern = oracle.synth_dense(kout * c, k * k, 42).reshape(kout, c, k, k)
    bias = oracle.synth_dense(kout, 1, 43).reshape(kout)
    hout, wout = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
    ref = oracle.transpose_mult_agg(oracle.image_flat(imgs, k, stride, pad, bx, by), oracle.kernel_flat(kern, bias, bx, by), 2)
    want = oracle.conv_result_to_images(ref, n, hout, wout, kout)
    got = np.loadtxt(tmp_path / "images.out.txt").reshape(n, kout, hout, wout)
    assert np.abs(got - want).max() <= TOL[prec] * np.abs(want).max()
    assert f"images={n}" in r.stdout
    assert f"result ( {n * hout * wout} X {kout} )" in r.stdout


@pytest.mark.parametrize("shape", [(2400, 9000, 700, 50, 500), (2048, 8200, 300, 20, 1000)])
def test_fused_layer1_kernel_vs_oracle(nb, oracle, shape):
    """Shapes large enough for the fused layer-1 kernel (fp64 -> fp16 conversion inside the persistent 2-SM GEMM through
    an L2-resident ring; whole 256-row tiles fused, left-over rows on the two-kernel path): labels and logits vs oracle."""
    batch, feat, hid, lab, blk = shape
    dev, ora, _ = _ff_sets(nb, oracle, batch, feat, hid, lab, blk)
    _, ref_logits, ref_labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], ora["x"], ora["b1"], ora["bo"], threads=4)
    for _ in range(2):  # twice: the ring and its flags are reused across calls
        dev["x"].invalidate_planes()
        _, labels, stats = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PREC_F16)
        assert np.array_equal(labels, ref_labels)
        z = nb.DeviceSet.lookup("ff", "__logits").download_dense()
        assert rel(z, ref_logits.dense()) <= TOL["f16"]


@pytest.mark.parametrize("prec", ["bf16x3", "fp64"])
def test_cpp_three_layer_inference(oracle, prec, tmp_path):
    """ff::inference (SimpleFF.cc:94-329): three separately executed graphs writing y1, y2, yo, then the softmax graph,
    each lowered to one block GEMM with its bias join fused; checked against the oracle's operators composed the same way."""
    from netsdb_b200 import build

    exe = build.build_host()["FFTest"]
    blk, batch, feat, hid, lab = 50, 130, 220, 110, 9
    dump = tmp_path / "dump3.txt"
    r = _run([exe, str(blk), str(blk), str(batch), str(feat), str(hid), str(lab), str(dump)],
             {"NB200_PRECISION": prec, "NB200_FF_LAYERS": "3"})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    got = np.loadtxt(dump)
    L = lambda rows, cols, seed, bc: oracle.load_matrix(oracle.synth_dense(rows, cols, seed), blk, bc)
    X, W1, B1 = L(batch, feat, 1, blk), L(hid, feat, 2, blk), L(hid, 1, 3, 1)
    WO, BO, W2, B2 = L(lab, hid, 4, blk), L(lab, 1, 5, 1), L(hid, hid, 6, blk), L(hid, 1, 7, 1)
    y1 = oracle.relu_bias_sum(oracle.transpose_mult_agg(W1, X, 2), B1)
    y2 = oracle.relu_bias_sum(oracle.input_layer_join_agg(W2, y1, 2), B2)
    z = oracle.input_layer_join_agg(WO, y2, 2)
    yo = oracle.transpose_bias_sum(z, BO, True)
    probs = oracle.output_layer(yo, oracle.row_aggregate(yo)).dense()
    assert np.isfinite(probs).all()
    assert np.array_equal(got[:, 1].astype(np.int32), np.argmax(probs, axis=1))   # argmax of the `output` set read back
    assert np.abs(got[:, 2] - probs[:, 0]).max() <= (1e-9 if prec == "fp64" else 1e-3)


def test_wide_hidden_layer_takes_two_kernel_path(nb, oracle):
    """hidden > 1,024 needs more than two 512-wide N tiles: the fused layer-1 kernel does not apply and the call must
    fall back to conversion + split-K GEMM (and still match the oracle)."""
    batch, feat, hid, lab, blk = 2100, 8300, 1100, 16, 500
    dev, ora, _ = _ff_sets(nb, oracle, batch, feat, hid, lab, blk)
    _, ref_logits, ref_labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], ora["x"], ora["b1"], ora["bo"], threads=4)
    _, labels, stats = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PREC_F16)
    assert np.array_equal(labels, ref_labels) and stats["splits1"] > 1
    assert rel(nb.DeviceSet.lookup("ff", "__logits").download_dense(), ref_logits.dense()) <= TOL["f16"]


@pytest.mark.parametrize("cfg", [(6, 10, 7, 3, 3, 2), (100, 500, 16, 1, 1, 4), (64, 300, 128, 10, 2, 3), (33, 77, 40, 5, 5, 1)])
def test_udf_networks_vs_oracle(nb, oracle, cfg):
    """nb200_udf_dense2 = FullyConnectedNetwork / SemanticClassifier (a two-layer network inside one SelectionComp) against
    the oracle's restatement, block by block; edge blocks of the input set are zero-padded as the reference's loader
    leaves them (the set is not a whole number of blocks in the last case)."""
    batch, feat, hid, lab, nbr, nbc = cfg
    rng = np.random.default_rng(batch + feat)
    rows, cols = nbr * batch - (3 if nbr > 2 else 0), nbc * feat
    x = rng.uniform(-1, 1, size=(rows, cols))
    w0, b0 = rng.uniform(-0.5, 0.5, size=(hid, feat)), rng.uniform(-0.5, 0.5, size=hid)
    w1, b1 = rng.uniform(-0.5, 0.5, size=(lab, hid)), rng.uniform(-0.5, 0.5, size=lab)
    ref = oracle.udf_dense2(oracle.load_matrix(x, batch, feat), feat, batch, w0, b0, w1, b1)
    d = nb.DeviceSet.create("udf", "x", rows, cols, batch, feat)
    d.upload_dense(x)
    got = nb.udf_dense2(d, feat, batch, w0, b0, w1, b1, "udf", "labels")
    assert got.info() == (nbr * lab, nbc * batch, lab, batch)
    dense = got.download_dense()
    n = 0
    for br, bc, _, _, arr in ref.blocks():
        assert np.array_equal(dense[br * lab:(br + 1) * lab, bc * batch:(bc + 1) * batch], arr), (br, bc)
        n += 1
    assert n == nbr * nbc and 0 < dense.mean() < 1
    with pytest.raises(nb.Nb200Error):   # a block with fewer than features x batch values: the UDF would read past it
        nb.udf_dense2(d, feat + 1, batch, np.zeros((hid, feat + 1)), b0, w1, b1, "udf", "bad")


def test_unmodified_reference_network_drivers():
    """netsDB's own FCProjTest.cc (FullyConnectedNetwork) and TestSemanticClassifier.cc (FFTransposeMult + FFAggMatrix +
    SemanticClassifier) compiled unmodified against the shim and run through executeComputations."""
    from netsdb_b200 import build

    fc, sc = os.path.join(build.HOST_BIN, "FCProjTest_netsdb"), os.path.join(build.HOST_BIN, "TestSemanticClassifier_netsdb")
    if not os.path.exists(fc) or not os.path.exists(sc):
        pytest.skip("reference drivers were not built (no reference tree at build time)")
    # the UDF insists on block rows == batch and views the block as features x batch: one 50 x 40 block (the driver ends
    # with sleep(20))
    r = _run([fc, "50", "40", "50", "40", "16", "2", "Y"], {"NB200_FC_SEED": "7"}, timeout=240)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert "End-to-End Time Duration" in r.stdout
    pairs = [ln for ln in r.stdout.splitlines() if ln.count(",") == 1 and set(ln.replace(" ", "")) <= set("01,")]
    assert len(pairs) >= 2      # the driver prints the first two values of each output row: labels, all 0 or 1
    r = _run([fc, "50", "40", "60", "40", "16", "2", "Y"], {"NB200_FC_SEED": "7"}, timeout=240)
    assert r.returncode != 0 and "batch size doesn't match" in (r.stdout + r.stderr)   # FullyConnectedNetwork.h:69-72


@pytest.mark.parametrize("rows", [9, 100])
def test_fp64_split_k_is_reproducible(nb, rows):
    """The fp64 path splits a long K over many CTAs; the partial tiles are summed in split order (k_f64_sum_splits), so a
    repeated call returns the same bits -- fp64 atomics would make the last bits depend on which CTA finished first."""
    K, N = 40000, 300
    a = nb.DeviceSet.create("rep", "a", rows, K, 1000, 1000); a.fill_synth(31)
    b = nb.DeviceSet.create("rep", "b", N, K, 1000, 1000); b.fill_synth(32)
    first = nb.gemm_blocks(nb.OP_NT, a, b, nb.EPI_NONE, None, nb.PREC_FP64, "rep", "c").download_dense()
    want = a.download_dense() @ b.download_dense().T
    assert np.abs(first - want).max() <= 1e-12 * np.abs(want).max()
    for _ in range(3):
        assert np.array_equal(nb.gemm_blocks(nb.OP_NT, a, b, nb.EPI_NONE, None, nb.PREC_FP64, "rep", "c").download_dense(), first)
