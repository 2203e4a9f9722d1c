"""Times the pieces of the label-exact cascade on the Amazon-14K shape in isolation (development probe):
the bf16x3 block GEMM of R undecided rows against W1 (operand planes cached), and the fp64 DMMA GEMM of a few rows."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from netsdb_b200 import capi as nb  # noqa: E402

nb.init(0)
F, H = 597540, 1000
w1 = nb.DeviceSet.create("p", "w1", H, F, 1000, 1000); w1.fill_synth(2)
for rows in (210, 105, 26, 9):
    x = nb.DeviceSet.create("p", "x", rows, F, 1000, 1000); x.fill_synth(1)
    for prec in ("bf16x3", "fp64") if rows <= 26 else ("bf16x3",):
        ts = []
        for rep in range(5):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n0 = nb.launch_count()
            e0.record()
            out = nb.gemm_blocks(nb.OP_NT, x, w1, nb.EPI_NONE, None, nb.PRECISIONS[prec], "p", "out")
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
            n1 = nb.launch_count() - n0
        print(f"rows={rows} {prec}: {min(ts[1:]):.3f} ms (launches {n1}, first call {ts[0]:.3f})", flush=True)
    x.destroy()
nb.shutdown()
