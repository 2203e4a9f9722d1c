"""CPU tests: the C oracle against the golden fixtures, numpy and its own invariants.

The oracle restates the reference's UDF arithmetic (oracle/ff_oracle.h cites file:line).  The reference ships no golden
vectors and its build cannot run here; these tests pin the oracle to hand-computed KATs, to an independent numpy
evaluation, and (test_reference_goldens) to vectors recorded from the reference's own UDF headers compiled unmodified
(oracle/ref); tests/test_oracle_pinned.py compares against that code directly where it was built."""
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(__file__), "golden")


def _load_ff(oracle, x, w1, b1, wo, bo, block):
    X = oracle.load_matrix(x, block, block)
    W1 = oracle.load_matrix(w1, block, block)
    B1 = oracle.load_matrix(b1, block, 1)       # FFTest.cc:109: (block_x, 1)
    WO = oracle.load_matrix(wo, block, block)   # FFTest.cc:112-113: (block_x, block_x)
    BO = oracle.load_matrix(bo, block, 1)
    return X, W1, B1, WO, BO


def test_generator_matches_numpy_restatement(oracle):
    a = oracle.synth_dense(13, 31, 42)
    assert np.array_equal(a, oracle.synth_dense_numpy(13, 31, 42))
    # distribution of ff::loadMatrix (FFMatrixUtil.cc:229-237): |v| in [1e-4, 0.5), both signs
    big = oracle.synth_dense(200, 500, 1)
    assert np.abs(big).min() >= 1e-4 and np.abs(big).max() < 0.5
    assert 0.45 < (big < 0).mean() < 0.55
    assert abs(np.abs(big).mean() - 0.25005) < 5e-3


def test_load_matrix_blocking_and_padding(oracle):
    d = np.arange(7 * 5, dtype=np.float64).reshape(7, 5) + 1
    s = oracle.load_matrix(d, 3, 2)
    blocks = {(br, bc): (tr, tc, arr.copy()) for br, bc, tr, tc, arr in s.blocks()}
    assert len(blocks) == 3 * 3
    assert all(v[2].shape == (3, 2) and v[0] == 7 and v[1] == 5 for v in blocks.values())
    assert np.array_equal(blocks[(2, 2)][2], [[35, 0], [0, 0], [0, 0]])  # zero padded edge block
    assert np.array_equal(s.dense(), d)
    t = oracle.load_matrix(d, 3, 2, dont_pad_rows=True, dont_pad_cols=True)
    shapes = {(br, bc): arr.shape for br, bc, _, _, arr in t.blocks()}
    assert shapes[(2, 2)] == (1, 1) and shapes[(0, 0)] == (3, 2)


def test_kat_2x3(oracle):
    g = np.load(os.path.join(G, "kat_2x3.npz"))
    X, W1, B1, WO, BO = _load_ff(oracle, g["x"], g["w1"], g["b1"], g["wo"], g["bo"], int(g["block"]))
    y1 = oracle.relu_bias_sum(oracle.transpose_mult_agg(W1, X, 2), B1)
    assert y1.totals() == (2, 2)                      # hidden x batch (FFTransposeMult.h:68-71)
    assert np.allclose(y1.dense(), g["y1"], atol=0, rtol=0)
    out, logits, labels = oracle.ff_inference_unit(W1, WO, X, B1, BO, threads=2)
    assert np.allclose(logits.dense(), g["logits"], atol=1e-15)
    assert np.allclose(out.dense(), g["probs"], atol=1e-15)
    assert np.array_equal(labels, g["labels"])


def test_ff_small_vs_numpy_golden(oracle):
    g = np.load(os.path.join(G, "ff_small.npz"))
    batch, feat, hid, lab, block = (int(g[k]) for k in ("batch", "feat", "hid", "lab", "block"))
    x = oracle.synth_dense(batch, feat, 1)
    w1 = oracle.synth_dense(hid, feat, 2)
    assert np.array_equal(x[:2, :5], g["x_head"]) and np.array_equal(w1[:2, :5], g["w1_head"])
    b1 = oracle.synth_dense(hid, 1, 3)
    wo = oracle.synth_dense(lab, hid, 4) * float(g["wo_scale"])
    bo = oracle.synth_dense(lab, 1, 5)
    for threads in (1, 3):
        out, logits, labels = oracle.ff_inference_unit(*[_load_ff(oracle, x, w1, b1, wo, bo, block)[i] for i in (1, 3, 0, 2, 4)],
                                                       threads=threads)
        assert out.totals() == (batch, lab)
        assert np.abs(logits.dense() - g["logits"]).max() <= 1e-12 * np.abs(g["logits"]).max()
        assert np.abs(out.dense() - g["probs"]).max() <= 1e-12
        assert np.array_equal(labels, g["labels"])


def test_output_block_meta_and_padding(oracle):
    """yo / output blocks: (in.blockCol, in.blockRow), J x I, totals swapped (FFTransposeBiasSum.h:64-67);
    exp only inside the un-padded region (:94-104) so padded entries stay 0 and do not enter the row sum."""
    batch, feat, hid, lab, block = 5, 6, 4, 3, 4
    x, w1, b1 = oracle.synth_dense(batch, feat, 1), oracle.synth_dense(hid, feat, 2), oracle.synth_dense(hid, 1, 3)
    wo, bo = oracle.synth_dense(lab, hid, 4), oracle.synth_dense(lab, 1, 5)
    X, W1, B1, WO, BO = _load_ff(oracle, x, w1, b1, wo, bo, block)
    z = oracle.input_layer_join_agg(WO, oracle.relu_bias_sum(oracle.transpose_mult_agg(W1, X), B1))
    assert z.totals() == (lab, batch)
    yo = oracle.transpose_bias_sum(z, BO)
    assert yo.totals() == (batch, lab)
    for br, bc, tr, tc, arr in yo.blocks():
        assert arr.shape == (block, block)
        rr, cc = np.meshgrid(br * block + np.arange(block), bc * block + np.arange(block), indexing="ij")
        valid = (rr < batch) & (cc < lab)
        label_pad = (rr < batch) & (cc >= lab)
        # padded label columns of real rows are exactly 0 (zero weights, zero bias, no exp); padded batch
        # rows carry relu(b1)-propagated values in the reference too and are don't-care (beyond totalRows)
        assert np.all(arr[label_pad] == 0.0) and np.all(arr[valid] > 0.0)
    rs = oracle.row_aggregate(yo)
    assert rs.totals() == (batch, 1)
    probs = oracle.output_layer(yo, rs).dense()
    assert np.allclose(probs.sum(axis=1), 1.0, atol=1e-14)


def test_join_dimension_mismatch_is_an_error(oracle):
    a = oracle.load_matrix(np.ones((4, 6)), 2, 3)
    b = oracle.load_matrix(np.ones((4, 6)), 2, 2)  # block K differs: "Block dimemsions mismatch!" FFTransposeMult.h:61-65
    with pytest.raises(RuntimeError):
        oracle.transpose_mult_agg(a, b)


def test_blocked_gemm_ops_vs_numpy(oracle):
    rng = np.random.default_rng(0)
    a, b = rng.standard_normal((23, 41)), rng.standard_normal((17, 41))
    for threads in (1, 4):
        c = oracle.transpose_mult_agg(oracle.load_matrix(a, 5, 7), oracle.load_matrix(b, 6, 7), threads)
        assert c.totals() == (23, 17)
        assert np.allclose(c.dense(), a @ b.T, rtol=0, atol=1e-12)
    b2 = rng.standard_normal((41, 9))
    c = oracle.input_layer_join_agg(oracle.load_matrix(a, 5, 7), oracle.load_matrix(b2, 7, 4), 2)
    assert c.totals() == (23, 9) and np.allclose(c.dense(), a @ b2, rtol=0, atol=1e-12)
    # empty input: loadMatrix returns without creating blocks when a dimension is 0 (FFMatrixUtil.cc:221-225)
    assert len(oracle.load_matrix(np.zeros((0, 5)), 2, 2)) == 0


def test_conv2d_memory_fusion_golden(oracle):
    g = np.load(os.path.join(G, "conv_small.npz"))
    IF = oracle.image_flat(g["images"], 3, 1, 0, 4, 5)
    assert np.array_equal(IF.dense(), g["image_flat"])
    assert IF.totals() == (2 * 4 * 3, 3 * 9 + 1)
    KF = oracle.kernel_flat(g["kernels"], g["bias"], 4, 5)
    assert np.array_equal(KF.dense(), g["kernel_flat"])
    res = oracle.transpose_mult_agg(IF, KF, 2)  # PipelinedConv2dMemFuseTest.cc:249-262
    assert np.allclose(res.dense(), g["result"], rtol=0, atol=1e-12)
    # padding > 0 and stride > 1 (Image.h:26-57)
    for k, s, p in ((3, 2, 1), (2, 1, 2)):
        assert np.array_equal(oracle.image_flat(g["images"], k, s, p, 8, 8).dense(), oracle.im2col_numpy(g["images"], k, s, p))


def test_word2vec_shape(oracle):
    """Word2Vec.cc:53-71: weights (dim x vocab) joined with inputs (batch x vocab) by FFTransposeMult."""
    w, x = oracle.synth_dense(6, 50, 9), oracle.synth_dense(11, 50, 10)
    out = oracle.transpose_mult_agg(oracle.load_matrix(w, 4, 16), oracle.load_matrix(x, 4, 16), 3)
    assert out.totals() == (6, 11) and np.allclose(out.dense(), w @ x.T, atol=1e-13)


def test_conv_result_to_images(oracle):
    """ConvResultToChunks + ConvChunksToImage restated chunk by chunk == the plain (image, window, channel) ->
    (image, channel, window) transpose, for block shapes that split images and channels unevenly."""
    rng = np.random.default_rng(1)
    n, z, x, y = 3, 5, 4, 6
    res = rng.normal(size=(n * x * y, z))
    want = res.reshape(n, x * y, z).transpose(0, 2, 1).reshape(n, z, x, y)
    for br, bc in ((7, 2), (24, 5), (10, 3), (100, 100), (1, 1)):
        assert np.array_equal(oracle.conv_result_to_images(oracle.load_matrix(res, br, bc), n, x, y, z), want)
    wide = oracle.conv_result_to_images(oracle.load_matrix(res, 7, 2), n, x, y, z + 2)  # more channels than columns
    assert np.array_equal(wide[:, :z], want) and not wide[:, z:].any()
    with pytest.raises(RuntimeError):
        oracle.conv_result_to_images(oracle.load_matrix(res, 7, 2), 5, x, y, z)  # 72 rows do not split into 5 images


def test_reference_goldens(oracle):
    """tests/golden/ref_pinned.npz was recorded from the REFERENCE'S OWN UDF code (oracle/_ref, netsDB's headers compiled
    unmodified; tests/golden/make_golden_ref.py).  The restatement must reproduce every vector: exactly where no product
    is involved, within summation-order noise (1e-13 of the scale) otherwise."""
    g = np.load(os.path.join(G, "ref_pinned.npz"))
    S = lambda m, br=10, bc=10: oracle.load_matrix(np.ascontiguousarray(m), br, bc)
    close = lambda a, b: np.abs(a - b).max() <= 1e-13 * max(np.abs(b).max(), 1.0)
    y1pre = oracle.transpose_mult_agg(S(g["ff_w1"]), S(g["ff_x"]), 2)
    assert close(y1pre.dense(), g["ff_y1pre"])
    y1 = oracle.relu_bias_sum(y1pre, S(g["ff_b1"], 10, 1))
    assert close(y1.dense(), g["ff_y1"])
    zpre = oracle.input_layer_join_agg(S(g["ff_wo"]), y1, 2)
    assert close(zpre.dense(), g["ff_zpre"])
    assert close(oracle.transpose_bias_sum_as_written(zpre, S(g["ff_bo"], 10, 1)).dense(), g["ff_yo_as_written"])
    assert close(oracle.row_aggregate(S(np.abs(g["ff_x"]))).dense(), g["ff_rowsum"])
    wx = oracle.input_layer_join_agg(S(g["lstm_w"]), S(g["lstm_xt"]), 2)
    uh = oracle.input_layer_join_agg(S(g["lstm_u"]), S(g["lstm_h"]), 2)
    assert close(oracle.lstm_sum(wx, uh, S(g["lstm_b"]), oracle.SIGMOID).dense(), g["lstm_gate_sigmoid"])
    assert close(oracle.lstm_sum(wx, uh, S(g["lstm_b"]), oracle.TANH).dense(), g["lstm_gate_tanh"])
    assert close(oracle.lstm_sum(wx, uh).dense(), g["lstm_two_sum"])
    assert close(oracle.lstm_hidden_state_agg(S(g["lstm_w"][:10, :20]), S(g["lstm_c"]), 2).dense(), g["lstm_hidden"])
    assert close(oracle.la_transpose_multiply_agg(S(g["la_a"]), S(g["la_a"]), 2).dense(), g["la_gram"])
    assert close(oracle.la_transpose_multiply_agg(S(g["la_a"]), S(g["la_b"]), 2).dense(), g["la_tn"])
    assert close(oracle.input_layer_join_agg(S(g["la_a"].T.copy()), S(g["la_b"]), 2).dense(), g["la_nn"])
    imf = oracle.image_flat(g["conv_imgs"], 3, 1, 0, 7, 5)
    assert np.array_equal(imf.dense(), g["conv_image_flat"])
    kf = oracle.kernel_flat(g["conv_kern"], np.zeros(5), 7, 5)
    assert np.array_equal(kf.dense(), g["conv_kernel_flat_nobias"])
    assert close(oracle.transpose_mult_agg(imf, kf, 2).dense(), g["conv_result_nobias"])
    assert np.array_equal(oracle.image_flat(g["conv_imgs2"], 3, 2, 1, 16, 4).dense(), g["conv_image_flat2"])
    assert np.array_equal(oracle.conv_result_to_images(S(g["conv_one_result"], 7, 2), 1, 4, 6, 5), g["conv_one_images"])
    stack = lambda s: np.stack([b[4] for b in sorted(s.blocks(), key=lambda b: (b[0], b[1]))])
    fc = oracle.udf_dense2(S(g["fc_x"], 6, 10), 10, 6, g["fc_w0"], g["fc_b0"], g["fc_w1"], g["fc_b1"], with_totals=False)
    assert np.array_equal(stack(fc), g["fc_out"])                  # FullyConnectedNetwork, the weights the reference drew
    sc = oracle.udf_dense2(S(g["sc_x"], 12, 5), 12, 5, np.ones((4, 12)), np.ones(4), np.ones((2, 4)), np.ones(2))
    assert np.array_equal(stack(sc), g["sc_out"])                  # SemanticClassifier
