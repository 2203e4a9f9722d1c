"""Generates tests/golden/*.npz.

The reference ships no golden vectors for the FF / conv2d / word2vec path, and
cannot be built or imported here, so these fixtures are NOT outputs of the reference: they are
(a) hand-computable known-answer cases written out explicitly below and (b) small seeded cases
evaluated by an independent dense numpy fp64 restatement (oracle/ff_oracle.py: ff_dense_numpy,
im2col_numpy).  They pin the C oracle and the CUDA path to each other and to numpy.

Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ff_oracle as fo  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def kat_2x3():
    """Hand-computable: batch 2, features 3, hidden 2, labels 2, blocks 2x2 (so every matrix is ragged)."""
    x = np.array([[1.0, 2.0, -1.0], [0.5, -1.0, 2.0]])
    w1 = np.array([[1.0, 0.0, 1.0], [-1.0, 1.0, 0.5]])
    b1 = np.array([[0.5], [-1.0]])
    wo = np.array([[1.0, -1.0], [0.5, 2.0]])
    bo = np.array([[0.0], [1.0]])
    # y1 = relu(w1 x^T + b1): w1 x^T = [[0, 2.5], [0.5, -0.5]] ; + b1 -> [[0.5, 3.0], [-0.5, -1.5]] -> relu [[0.5, 3.0], [0, 0]]
    y1 = np.array([[0.5, 3.0], [0.0, 0.0]])
    # z^T = wo y1 + bo = [[0.5, 3.0], [0.25, 1.5]] + [[0],[1]] = [[0.5, 3.0], [1.25, 2.5]] ; z = batch x labels
    z = np.array([[0.5, 1.25], [3.0, 2.5]])
    e = np.exp(z)
    probs = e / e.sum(axis=1, keepdims=True)
    labels = np.array([1, 0], dtype=np.int32)
    return dict(x=x, w1=w1, b1=b1, wo=wo, bo=bo, y1=y1, logits=z, probs=probs, labels=labels, block=np.int32(2))


def seeded_ff(batch, feat, hid, lab, block, wo_scale):
    x = fo.synth_dense_numpy(batch, feat, 1)
    w1 = fo.synth_dense_numpy(hid, feat, 2)
    b1 = fo.synth_dense_numpy(hid, 1, 3)
    wo = fo.synth_dense_numpy(lab, hid, 4) * wo_scale
    bo = fo.synth_dense_numpy(lab, 1, 5)
    probs, z, labels = fo.ff_dense_numpy(x, w1, b1, wo, bo)
    return dict(batch=np.int32(batch), feat=np.int32(feat), hid=np.int32(hid), lab=np.int32(lab), block=np.int32(block),
                wo_scale=np.float64(wo_scale), logits=z, probs=probs, labels=labels,
                x_head=x[:2, :5].copy(), w1_head=w1[:2, :5].copy())


def conv_case():
    rng = np.random.default_rng(7)
    imgs = rng.standard_normal((2, 3, 6, 5))
    kern = rng.standard_normal((4, 3, 3, 3))
    bias = rng.standard_normal(4)
    flat = fo.im2col_numpy(imgs, 3, 1, 0)
    kflat = np.concatenate([kern.reshape(4, -1), bias.reshape(4, 1)], axis=1)
    result = flat @ kflat.T  # FFTransposeMult + FFAggMatrix on (image_flat, kernel_flat)
    # independent check of the im2col-as-GEMM identity against a direct convolution
    direct = np.zeros((2, 4, 4, 3))
    for n in range(2):
        for o in range(4):
            for r in range(4):
                for c in range(3):
                    direct[n, o, r, c] = (imgs[n, :, r:r + 3, c:c + 3] * kern[o]).sum() + bias[o]
    assert np.allclose(result.reshape(2, 4, 3, 4).transpose(0, 3, 1, 2), direct)  # (n, ho, wo, cout) -> (n, cout, ho, wo)
    return dict(images=imgs, kernels=kern, bias=bias, image_flat=flat, kernel_flat=kflat, result=result)


if __name__ == "__main__":
    np.savez(os.path.join(HERE, "kat_2x3.npz"), **kat_2x3())
    np.savez(os.path.join(HERE, "ff_small.npz"), **seeded_ff(37, 53, 29, 7, 10, 0.01))
    np.savez(os.path.join(HERE, "conv_small.npz"), **conv_case())
    print("golden fixtures written to", HERE)
