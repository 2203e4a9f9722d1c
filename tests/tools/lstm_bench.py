"""Times one LSTM step (netsdb_b200.lstm.lstm_cell: the graph of src/tests/source/LSTMTest.cc) with all sets resident
on the device, and the CPU oracle on the same sizes.  Device time by CUDA events on the library's stream."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from netsdb_b200 import capi, lstm

L, D, B, bs = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (2048, 4096, 2048, 1024)
capi.init(0)
torch.cuda.set_device(0)
stream = torch.cuda.Stream()
capi.set_stream(stream.cuda_stream)
shapes = {"x_t": (D, B), "h_t_1": (L, B), "c_t_1": (L, B)}
for g in "fioc":
    shapes.update({"w_" + g: (L, D), "u_" + g: (L, L), "b_" + g: (L, B)})
for i, (k, (r, c)) in enumerate(sorted(shapes.items())):
    capi.DeviceSet.create("lstm", k, r, c, bs, bs).fill_synth(100 + i)
flops = 4 * (2 * L * D * B + 2 * L * L * B) + 3 * 2 * L * min(L, B) * B
for prec in ("bf16", "bf16x3", "composed:f16", "composed:bf16x3", "composed:fp64"):
    fn = lstm.lstm_cell if prec.startswith("composed:") else lstm.lstm_cell_fused
    pname = prec.split(":")[-1]
    times = []
    for it in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            n0 = capi.launch_count()
            fn("lstm", pname)
            n1 = capi.launch_count()
            e1.record(stream)
        e1.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = float(np.median(times[2:]))
    print(f"lstm cell L={L} D={D} B={B} {prec}: {ms:.3f} ms/step, {flops / ms / 1e9:.1f} useful TFLOP/s, "
          f"{n1 - n0} kernel launches", flush=True)
if "--cpu" in sys.argv:
    from oracle import ff_oracle as fo
    p = {k: fo.load_matrix_synth(100 + i, r, c, bs, bs) for i, (k, (r, c)) in enumerate(sorted(shapes.items()))}
    t0 = time.perf_counter()
    fo.lstm_cell(p, threads=fo.hw_threads())
    dt = time.perf_counter() - t0
    print(f"oracle (CPU port, {fo.hw_threads()} threads): {dt * 1e3:.1f} ms/step, {flops / dt / 1e9:.1f} GFLOP/s", flush=True)
capi.shutdown()
