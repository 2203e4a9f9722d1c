"""One process, several GPUs (what a netsDB backend with `numThreads` pipeline threads needs, Configuration.h:52-53,
PDBWorkerQueue.cc:163): every thread picks its device with nb200_use_device, holds its row shard of `inputs` and the
replicated weights there, and runs the fused stage; the labels of the shards equal the 1-GPU run of the whole batch.
Needs two visible GPUs (skipped on a 1-GPU box; run under `gpurun --gpus 2`)."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_two_devices_from_two_threads_of_one_process(nb):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from netsdb_b200.shard import row_shard

    batch, feat, hid, lab, blk = 2500, 8320, 1000, 64, 1000
    nb.init_devices([0, 1])

    def sets(rows, row0):
        DS = nb.DeviceSet
        w1 = DS.create("md", "w1", hid, feat, blk, blk); w1.fill_synth(2)
        b1 = DS.create("md", "b1", hid, 1, blk, 1); b1.fill_synth(3)
        wo = DS.create("md", "wo", lab, hid, blk, blk); wo.fill_synth(4)
        bo = DS.create("md", "bo", lab, 1, blk, 1); bo.fill_synth(5)
        x = DS.create("md", "inputs", rows, feat, blk, blk); x.fill_synth(1, row_offset=row0)
        return w1, wo, x, b1, bo

    nb.use_device(0)
    full = sets(batch, 0)
    _, ref_labels, _ = nb.ff_inference_unit(*full, nb.PREC_F16, "md", "out_full")
    got, errs, foreign = {}, [], {}

    def work(d):
        try:
            nb.use_device(d)
            row0, rows = row_shard(batch, 2, d)
            s = sets(rows, row0)                       # the (db, set) namespace is per device: same names, other device
            foreign[d] = s[2]
            for _ in range(3):
                s[2].invalidate_planes()
                _, labels, st = nb.ff_inference_unit(*s, nb.PREC_F16, "md", "out")
            got[d] = labels
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=work, args=(d,)) for d in (0, 1)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    assert np.array_equal(np.concatenate([got[0], got[1]]), ref_labels)
    # a set of device 1 handed to a thread that works on device 0 is refused, not silently dereferenced on the wrong GPU
    nb.use_device(0)
    with pytest.raises(nb.Nb200Error):
        nb.ff_inference_unit(full[0], full[1], foreign[1], full[3], full[4], nb.PREC_F16, "md", "bad")
    nb.use_device(0)


def test_cpp_driver_shards_inputs_over_two_devices(oracle, tmp_path):
    """The C++ host shim with NB200_DEVICES=2: PDBClient::sendData splits the rows of `inputs` over the devices (row
    granular: a block can straddle two devices) and replicates the other sets, executeComputations lowers the graph once
    per device on its own thread, getSetIterator assembles the blocks of `output` from the devices that hold their rows.
    The same FFTest driver, the same oracle check as on one device."""
    import os
    import subprocess

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from netsdb_b200 import build

    exe = build.build_host()["FFTest"]
    blk, batch, feat, hid, lab = 100, 250, 640, 300, 12     # 250 rows / 2 = 125: block row 1 straddles the two devices
    ref = None
    for ndev in (2, 1):
        dump = tmp_path / f"dump{ndev}.txt"
        env = dict(os.environ, NB200_PRECISION="f16", NB200_DEVICES=str(ndev))
        r = subprocess.run([exe, str(blk), str(blk), str(batch), str(feat), str(hid), str(lab), str(dump)], env=env,
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        got = np.loadtxt(dump)
        if ref is None:
            X = oracle.load_matrix(oracle.synth_dense(batch, feat, 1), blk, blk)
            W1 = oracle.load_matrix(oracle.synth_dense(hid, feat, 2), blk, blk)
            B1 = oracle.load_matrix(oracle.synth_dense(hid, 1, 3), blk, 1)
            WO = oracle.load_matrix(oracle.synth_dense(lab, hid, 4), blk, blk)
            BO = oracle.load_matrix(oracle.synth_dense(lab, 1, 5), blk, 1)
            ref_out, _, ref_labels = oracle.ff_inference_unit(W1, WO, X, B1, BO, threads=2)
            ref = (ref_out.dense(), ref_labels)
        assert np.array_equal(got[:, 0].astype(np.int32), ref[1]), ndev      # fused stage's labels, concatenated over devices
        assert np.array_equal(got[:, 1].astype(np.int32), ref[1]), ndev      # argmax of the `output` set read back
        assert np.abs(got[:, 2] - ref[0][:, 0]).max() <= 1e-2


def test_network_udf_over_a_sharded_set_is_refused():
    """FullyConnectedNetwork reinterprets whole blocks; with NB200_DEVICES=2 the rows of `inputs` are split over the devices
    at row granularity, so the lowering refuses instead of evaluating half-empty blocks.  With the set taken out of
    NB200_SHARD_SETS the same driver runs."""
    import os
    import subprocess

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from netsdb_b200 import build

    exe = os.path.join(build.HOST_BIN, "FCProjTest_netsdb")
    if not os.path.exists(exe):
        pytest.skip("reference driver was not built (no reference tree at build time)")
    args = [exe, "50", "40", "50", "40", "16", "2", "Y"]
    r = subprocess.run(args, env=dict(os.environ, NB200_DEVICES="2", NB200_FC_SEED="7"), capture_output=True, text=True, timeout=240)
    assert r.returncode != 0 and "sharded by NB200_DEVICES" in (r.stdout + r.stderr)
    r = subprocess.run(args, env=dict(os.environ, NB200_DEVICES="2", NB200_FC_SEED="7", NB200_SHARD_SETS="none"), capture_output=True,
                       text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
