"""GPU parity tests of the streamed layer-1 kernel (netsdb_b200/csrc/nb200_l1stream.cu): FFTransposeMult + FFAggMatrix over
the fp64 `inputs` blocks as they lie in HBM (FFTransposeMult.h:80-92, CombinedShuffleSink.h:112-144), through the C ABI,
against the CPU oracle and fp64 numpy.  Shapes: the 1,250-row shard of BASELINE config 3 at 8 GPUs, a ragged K tail, a
batch that is not a multiple of any tile, a handful of rows; both cluster shapes (4 and 8 CTAs)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL_F16 = 2.5e-3   # of max|z|, the bound of the single-pass fp16 variant (tests/test_gpu_parity.py)


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


@pytest.fixture
def stream_opts(nb):
    yield nb
    nb.set_option("fused_layer1", -1)
    nb.set_option("l1_stream_mt", 0)


def _sets(nb, oracle, batch, feat, hid, lab, blk):
    dense = dict(x=oracle.synth_dense(batch, feat, 1), w1=oracle.synth_dense(hid, feat, 2), b1=oracle.synth_dense(hid, 1, 3),
                 wo=oracle.synth_dense(lab, hid, 4), bo=oracle.synth_dense(lab, 1, 5))
    dims = dict(x=(batch, feat, blk, blk), w1=(hid, feat, blk, blk), b1=(hid, 1, blk, 1), wo=(lab, hid, blk, blk), bo=(lab, 1, blk, 1))
    names = dict(x="inputs", w1="w1", b1="b1", wo="wo", bo="bo")
    dev, ora = {}, {}
    for k, (r, c, br, bc) in dims.items():
        dev[k] = nb.DeviceSet.create("l1s", names[k], r, c, br, bc)
        dev[k].upload_dense(dense[k])
        ora[k] = oracle.load_matrix(dense[k], br, bc)
    return dev, ora, dense


@pytest.mark.parametrize("mt", [1, 2])
@pytest.mark.parametrize("shape", [(1250, 8256, 1000, 40, 1000), (300, 8228, 1000, 17, 500), (2600, 4160, 1024, 9, 520), (5, 4100, 600, 7, 100),
                                   (900, 8200, 500, 12, 1000), (640, 4608, 300, 33, 512)])
def test_inference_unit_through_the_stream_kernel_vs_oracle(stream_opts, oracle, shape, mt):
    nb = stream_opts
    batch, feat, hid, lab, blk = shape
    dev, ora, _ = _sets(nb, oracle, batch, feat, hid, lab, blk)
    _, ref_logits, ref_labels = oracle.ff_inference_unit(ora["w1"], ora["wo"], ora["x"], ora["b1"], ora["bo"], threads=4)
    nb.set_option("fused_layer1", 2)
    nb.set_option("l1_stream_mt", mt)
    for _ in range(2):  # twice: the wave plan and the workspace are reused across calls
        dev["x"].invalidate_planes()
        _, labels, stats = nb.ff_inference_unit(dev["w1"], dev["wo"], dev["x"], dev["b1"], dev["bo"], nb.PREC_F16)
        assert np.array_equal(labels, ref_labels), f"{(labels != ref_labels).sum()} labels differ"
        assert rel(nb.DeviceSet.lookup("ff", "__logits").download_dense(), ref_logits.dense()) <= TOL_F16
        assert stats["splits1"] >= 2   # few tiles, many clusters: K is split (the FFAggMatrix sum over partial tiles)


@pytest.mark.parametrize("mt", [1, 2])
def test_layer1_partial_three_ways(stream_opts, mt):
    """nb200_ff_layer1_partial (the pre-bias sums of layer 1 and the per-row sums of squares): the stream kernel, the ring
    kernel's fall-back and the separate kernels against fp64 numpy, and against one another (same fp16 operands, so they
    differ in summation order only)."""
    import torch

    nb = stream_opts
    rows, F, H = 700, 12000, 777
    w1 = nb.DeviceSet.create("l1s", "w1", H, F, 1000, 1000); w1.fill_synth(2)
    x = nb.DeviceSet.create("l1s", "x", rows, F, 1000, 1000); x.fill_synth(1)
    X, W = x.download_dense(), w1.download_dense()
    want, ssq = X @ W.T, (X ** 2).sum(axis=1)
    got = {}
    nb.set_option("l1_stream_mt", mt)
    for mode in (2, 0):
        nb.set_option("fused_layer1", mode)
        x.invalidate_planes()
        out = torch.zeros((rows, H + 1), dtype=torch.float32, device="cuda")
        n0 = nb.launch_count()
        nb.check(nb.lib().nb200_ff_layer1_partial(w1.ptr, x.ptr, nb.PREC_F16, C.c_void_p(out.data_ptr()), H + 1))
        nb.synchronize()
        torch.cuda.synchronize()
        got[mode] = out.cpu().numpy()
        assert nb.launch_count() > n0
        assert rel(got[mode][:, :H], want) <= TOL_F16
        assert np.abs(got[mode][:, H] - ssq).max() <= 1e-4 * ssq.max()
    assert rel(got[2][:, :H], got[0][:, :H]) <= 1e-5
