"""Records golden vectors from the REFERENCE'S OWN UDF code (oracle/_ref/libff_ref.so = netsDB's UDF headers compiled
unmodified, see oracle/ref/ref_harness.cc) into tests/golden/ref_pinned.npz.  Run where /root/reference is mounted:

    python tests/golden/make_golden_ref.py

Every entry is a dense input or the dense (padding stripped) output the reference produced for it; the tests replay the
inputs through the oracle (tests/test_oracle.py::test_reference_goldens) and through the device
(tests/test_gpu_parity.py::test_device_against_reference_goldens)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ff_oracle as fo, ff_ref as fr  # noqa: E402

assert fr.build(), "needs the reference tree"
rng = np.random.default_rng(20240607)
g = {}
BS = 10

# the FF pipeline, operator by operator (batch 23, features 37, hidden 19, labels 12, blocks of 10)
x, w1, b1 = rng.normal(size=(23, 37)), rng.normal(size=(19, 37)) * 0.3, rng.normal(size=(19, 1))
wo, bo = rng.normal(size=(12, 19)) * 0.3, rng.normal(size=(12, 1))
S = lambda m, br=BS, bc=BS: fo.load_matrix(m, br, bc)
y1pre = fr.transpose_mult_agg(S(w1), S(x))
y1 = fr.relu_bias_sum(y1pre, S(b1, BS, 1))
zpre = fr.input_layer_join_agg(S(wo), y1)
g.update(ff_x=x, ff_w1=w1, ff_b1=b1, ff_wo=wo, ff_bo=bo, ff_y1pre=y1pre.dense(), ff_y1=y1.dense(), ff_zpre=zpre.dense(),
         ff_yo_as_written=fr.transpose_bias_sum_as_written(zpre, S(bo, BS, 1)).dense(),
         ff_rowsum=fr.row_aggregate(S(np.abs(x))).dense())

# LSTM operators (L = B = 20, D = 30)
L, D, B = 20, 30, 20
for k, shape in dict(w=(L, D), u=(L, L), b=(L, B), xt=(D, B), h=(L, B), c=(L, B)).items():
    g["lstm_" + k] = rng.uniform(-0.5, 0.5, size=shape)
wx = fr.input_layer_join_agg(S(g["lstm_w"]), S(g["lstm_xt"]))
uh = fr.input_layer_join_agg(S(g["lstm_u"]), S(g["lstm_h"]))
g["lstm_gate_sigmoid"] = fr.lstm_sum(wx, uh, S(g["lstm_b"]), 1).dense()
g["lstm_gate_tanh"] = fr.lstm_sum(wx, uh, S(g["lstm_b"]), 2).dense()
g["lstm_two_sum"] = fr.lstm_sum(wx, uh).dense()
g["lstm_hidden"] = fr.lstm_hidden_state_agg(S(g["lstm_w"][:10, :20]), S(g["lstm_c"])).dense()  # one row block: tanh once

# LA DSL
la_a, la_b = rng.normal(size=(40, 20)), rng.normal(size=(40, 17))
g.update(la_a=la_a, la_b=la_b, la_gram=fr.la_multiply_agg(S(la_a), S(la_a), transpose_left=True).dense(),
         la_tn=fr.la_multiply_agg(S(la_a), S(la_b), transpose_left=True).dense(),
         la_nn=fr.la_multiply_agg(S(la_a.T.copy()), S(la_b)).dense())

# conv2d_memory_fusion (3 images of 2 x 9 x 8, 5 kernels of 3 x 3, stride 1, no padding; blocks 7 x 5)
imgs, kern, bias = rng.normal(size=(3, 2, 9, 8)), rng.normal(size=(5, 2, 3, 3)), rng.normal(size=5)
image_flat = fr.image_flat(imgs, 3, 1, 0, 7, 5)
kernel_noj = fr.kernel_flat(kern, None, 7, 5, with_join=False)
g.update(conv_imgs=imgs, conv_kern=kern, conv_bias=bias, conv_image_flat=image_flat.dense(), conv_kernel_flat_nobias=kernel_noj.dense(),
         conv_result_nobias=fr.transpose_mult_agg(image_flat, kernel_noj).dense())
imgs2 = rng.normal(size=(2, 3, 10, 10))
g.update(conv_imgs2=imgs2, conv_image_flat2=fr.image_flat(imgs2, 3, 2, 1, 16, 4).dense())
one = rng.normal(size=(24, 5))
g.update(conv_one_result=one, conv_one_images=fr.conv_result_to_images(S(one, 7, 2), 1, 4, 6, 5))

# the two networks that live inside one SelectionComp: FullyConnectedNetwork (weights drawn by the reference's own
# constructor, read back out of the object) and SemanticClassifier (all-ones weights)
fc_x = rng.uniform(-2, 2, size=(3 * 6, 2 * 10))                      # 3 x 2 blocks of batch 6 x features 10
fc_out, fc_w0, fc_b0, fc_w1, fc_b1 = fr.fully_connected(S(fc_x, 6, 10), 10, 3, 7, 6)
g.update(fc_x=fc_x, fc_w0=fc_w0, fc_b0=fc_b0, fc_w1=fc_w1, fc_b1=fc_b1,
         fc_out=np.stack([b[4] for b in sorted(fc_out.blocks(), key=lambda b: (b[0], b[1]))]))
sc_x = rng.uniform(-1, 0.6, size=(12, 3 * 5))                        # embedding 12, three blocks of batch 5
sc_out = fr.semantic_classifier(S(sc_x, 12, 5), 12, 4, 2)
g.update(sc_x=sc_x, sc_out=np.stack([b[4] for b in sorted(sc_out.blocks(), key=lambda b: (b[0], b[1]))]))

np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_pinned.npz"), **g)
print("wrote", len(g), "arrays")
