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
