"""bench.py --config c4 | c5: the other two BASELINE.json configurations, one JSON line each in bench.py's shape
(metric / value / roofline / cpu_baseline / e2e).  They are parity-test cases first (tests/test_gpu_parity.py,
tests/test_gpu_fullsize.py); these lines put a reproducible number beside them.

c4  conv2d_memory_fusion: 3x3 conv, 64 -> 64 channels, 112 x 112 input, batch 64, + ReLU (PipelinedConv2dMemFuseTest.cc:
    136-273: image_ops + kernel_bias_join + conv2d jobs; the ReLU is FFReluBiasSum's max(0, .) applied to the result)
c5  word2vec embedding inference: weights 500 x 1,000,000, batch 100,000 (Word2Vec.cc:45-98 at BASELINE's scale-up): the
    batch streams through one GPU in row slabs generated on the device; one step = one slab
"""
from __future__ import annotations

import json
import os
import time

ROOT = os.path.dirname(os.path.abspath(__file__))


def _peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["bf16_tflops"]), float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 1590.0, 6650.0, "fallback"


def run_c4(a):
    import numpy as np
    import torch

    from netsdb_b200 import capi

    torch.cuda.set_device(0)
    capi.init(0)
    stream = torch.cuda.current_stream()
    capi.set_stream(stream.cuda_stream)
    n, c, h, w, kout, k, pad = 64, 64, 112, 112, 64, 3, 0
    hout, wout = h - k + 1, w - k + 1
    rng = np.random.default_rng(0)
    imgs_t = torch.empty(n * c * h * w, dtype=torch.float64).pin_memory()
    imgs = imgs_t.numpy().reshape(n, c, h, w)
    imgs[:] = rng.uniform(-0.5, 0.5, size=(n, c, h, w))
    kern = rng.uniform(-0.5, 0.5, size=(kout, c, k, k))
    bias = rng.uniform(-0.5, 0.5, size=kout)
    prec = capi.PRECISIONS["bf16" if a.precision == "f16" else a.precision]
    flops = 2.0 * n * hout * wout * kout * c * k * k
    # device-resident: the image set uploaded once (its NHWC operand planes are derived once and cached with the set)
    iset = capi.conv2d_images_create(imgs, "conv2d", "img")
    for _ in range(max(a.warmup, 3)):
        R = capi.conv2d_fused_from_set(iset, kern, bias, pad, True, prec, 1000, 1000)
    torch.cuda.synchronize()
    l0 = capi.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kms = []
    e0.record(stream)
    for _ in range(a.steps):
        R = capi.conv2d_fused_from_set(iset, kern, bias, pad, True, prec, 1000, 1000)
        kms.append(capi.last_kernel_ms())
    e1.record(stream)
    torch.cuda.synchronize()
    launches = capi.launch_count() - l0
    ms = e0.elapsed_time(e1) / a.steps
    k_ms = float(np.median(kms))
    # e2e: host images (pinned fp64 NCHW, as the reference's Image records hold them) in, result blocks out, per step
    out_host = torch.empty(n * hout * wout * kout, dtype=torch.float64).pin_memory().numpy().reshape(n * hout * wout, kout)
    def e2e_step():
        Rr = capi.conv2d_fused(imgs, kern, bias, pad, True, prec, 1000, 1000)
        capi.check(capi.lib().nb200_download_rect(Rr.ptr, 0, 0, out_host.ctypes.data, n * hout * wout, kout, kout))
    e2e_step()
    t0 = time.perf_counter()
    for _ in range(a.e2e_steps):
        e2e_step()
    capi.synchronize()
    e_ms = (time.perf_counter() - t0) * 1e3 / a.e2e_steps
    # parity of the timed path on a sample of output positions (fp64 numpy of the same conv + ReLU)
    got = R.download_dense()
    idx = rng.integers(0, n * hout * wout, size=64)
    ref = np.empty((64, kout))
    for t, r in enumerate(idx):
        im, pos = divmod(int(r), hout * wout)
        oy, ox = divmod(pos, wout)
        ref[t] = np.maximum((kern * imgs[im, :, oy:oy + k, ox:ox + k][None]).sum(axis=(1, 2, 3)) + bias, 0.0)
    err = float(np.abs(got[idx] - ref).max() / np.abs(ref).max())
    peak_tf, peak_hbm, src = _peaks()
    alg_bytes = n * h * w * c * 2.0 + n * hout * wout * kout * 4.0   # bf16 NHWC image read once + fp32 result written once
    traffic = None   # DRAM bytes per launch from the committed ncu capture, while the kernel source is the one profiled
    try:
        import hashlib
        root = os.path.dirname(os.path.abspath(__file__))
        tj = json.load(open(os.path.join(root, "profiles", "r02_traffic_conv.json")))
        sha = hashlib.sha256(open(os.path.join(root, "netsdb_b200", "csrc", "nb200_gemm.cu"), "rb").read()).hexdigest()[:16]
        if tj.get("kernel_source_sha16") == sha:
            traffic = tj["k_conv_persist"]["dram_bytes_read"] + tj["k_conv_persist"]["dram_bytes_write"]
    except (OSError, KeyError, ValueError):
        pass
    cpu = None
    if not a.no_cpu_baseline:
        from oracle import ff_oracle as fo
        ns = 4                                          # images of the sample (the job is linear in the image count)
        t0 = time.perf_counter()
        flat = fo.image_flat(imgs[:ns], k, 1, pad, 1000, 1000)          # ImageToChunks + ImageChunksToBlock + ImageBlockToMatrix
        kf = fo.kernel_flat(kern, bias, 1000, 1000)                     # KernelToChunks + KernelBiasJoin
        fo.transpose_mult_agg(flat, kf, fo.hw_threads())                # FFTransposeMult + FFAggMatrix (the conv2d job)
        dt = time.perf_counter() - t0
        cpu = dict(value=ns / dt, unit="images/s", cores=fo.hw_threads(), kind="port", sample_seconds=round(dt, 3),
                   sample=f"{ns} of {n} images through the oracle's image_flat + kernel_flat + block GEMM jobs, fp64, ReLU omitted (negligible)")
    line = dict(metric="conv2d_memory_fusion images/s (3x3, 64->64, 112x112, batch 64, + ReLU)", value=n / (ms / 1e3), unit="images/s",
                n_gpus=1, steps=a.steps, warmup=max(a.warmup, 3), ms_per_step=ms, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="bf16" if prec == capi.PRECISIONS["bf16"] else a.precision, data="synthetic",
                config=dict(workload="conv2d_memory_fusion: 3x3 conv, 64->64 channels, 112x112 input, batch 64, stride 1, pad 0 (out 110x110), + ReLU",
                            l2="operands fit L2: every step reads the same image set (stated, not flushed); the result (198 MB) does not",
                            logit_tolerance=f"sampled outputs within {err:.1e} of max|result| of fp64 numpy"),
                e2e=dict(value=n / (e_ms / 1e3), unit="images/s", h2d_bytes_per_step=int(imgs.nbytes), d2h_bytes_per_step=int(out_host.nbytes),
                         ms_per_step=e_ms, note="pinned fp64 NCHW images in, fp64 result blocks out, every step"),
                gpu_launches=int(launches),
                roofline=dict(bound="hbm", kernel="k_conv_persist (persistent 4-D TMA implicit GEMM: filter planes resident, each input row loaded once, bias + ReLU + TMA-store epilogue)",
                              achieved=alg_bytes / (k_ms / 1e3) / 1e9,
                              peak=peak_hbm, unit="GB/s", frac=alg_bytes / (k_ms / 1e3) / 1e9 / peak_hbm, traffic=traffic, peak_source=src,
                              ms_per_launch=k_ms, tensor_tflops=flops / (k_ms / 1e3) / 1e12, tensor_frac=flops / (k_ms / 1e3) / 1e12 / peak_tf),
                cpu_baseline=cpu, parity_sample_rel_err=err)
    print(json.dumps(line), flush=True)
    capi.shutdown()


def run_c5(a):
    import ctypes as C

    import numpy as np
    import torch

    from netsdb_b200 import capi

    torch.cuda.set_device(0)
    capi.init(0)
    stream = torch.cuda.current_stream()
    capi.set_stream(stream.cuda_stream)
    dim, vocab, batch, blk = 500, 1000000, 100000, 1000
    slab = 10000                                    # rows resident at a time: 80 GB of fp64
    DS = capi.DeviceSet
    W = DS.create("word2vec", "weights0", dim, vocab, blk, blk); W.fill_synth(21)
    X = DS.create("word2vec", "inputs", slab, vocab, blk, blk)
    out = torch.zeros((slab, dim), dtype=torch.float32, device="cuda")
    prec = capi.PRECISIONS[a.precision]

    def step(i):
        X.fill_synth(22, row_offset=(i % (batch // slab)) * slab)      # the slab's rows of the 100,000 x 1,000,000 matrix (untimed)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        # FFTransposeMult + FFAggMatrix on (weights, inputs): pre-bias sums, batch x dim (Word2Vec.cc:61-77)
        capi.check(capi.lib().nb200_ff_layer1_partial(W.ptr, X.ptr, prec, C.c_void_p(out.data_ptr()), dim))
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    for i in range(max(a.warmup, 3)):
        step(i)
    l0 = capi.launch_count()
    ts = [step(i) for i in range(a.steps)]
    launches = capi.launch_count() - l0
    ms = float(np.mean(ts))
    # parity on a sample: rows of the last slab against fp64 numpy of the same (downloaded) operands
    ridx = np.array([0, 4321, slab - 1])
    xs = np.empty((len(ridx), vocab))
    for t, r in enumerate(ridx):
        capi.check(capi.lib().nb200_download_rect(X.ptr, int(r), 0, xs[t].ctypes.data, 1, vocab, vocab))
    wsub = np.empty((16, vocab))
    capi.check(capi.lib().nb200_download_rect(W.ptr, 0, 0, wsub.ctypes.data, 16, vocab, vocab))
    ref = xs @ wsub.T
    got = out.cpu().numpy()[ridx][:, :16]
    err = float(np.abs(got - ref).max() / np.abs(ref).max())
    peak_tf, peak_hbm, src = _peaks()
    flops = 2.0 * slab * vocab * dim
    alg_bytes = slab * vocab * 8.0 + dim * vocab * 2.0
    line = dict(metric="word2vec embedding inference rows/s (vocab 1M x dim 500, batch 100,000 streamed in 10,000-row slabs)",
                value=slab / (ms / 1e3), unit="rows/s", n_gpus=1, steps=a.steps, warmup=max(a.warmup, 3), ms_per_step=ms, higher_is_better=True,
                scaling="weak", vs_baseline=None, dtype=a.precision, data="synthetic",
                config=dict(workload="word2vec embedding inference: synthetic vocab 1M x dim 500 embedding block join over batch 100,000; one step = one 10,000-row slab (80 GB fp64) resident in HBM",
                            l2="the slab (80 GB) exceeds L2, no flush needed",
                            logit_tolerance=f"sampled outputs within {err:.1e} of max|result| of fp64 numpy (bound 2.5e-3 for f16)"),
                e2e=None, gpu_launches=int(launches),
                roofline=dict(bound="hbm", kernel="nb200_ff_layer1_partial: k_l1_stream (fp64 inputs streamed once, converted in registers, tcgen05 cta_group::2; the hidden half past dim 500 multiplies zero rows) + k_l1s_reduce", achieved=alg_bytes / (ms / 1e3) / 1e9,
                              peak=peak_hbm, unit="GB/s", frac=alg_bytes / (ms / 1e3) / 1e9 / peak_hbm, traffic=None, peak_source=src,
                              tensor_tflops=flops / (ms / 1e3) / 1e12, tensor_frac=flops / (ms / 1e3) / 1e12 / peak_tf),
                cpu_baseline=None, parity_sample_rel_err=err,
                note="e2e omitted: 800 GB of fp64 inputs per batch do not exist on the host of this box; the reference's own size is batch 100 (tests/test_gpu_fullsize.py)")
    print(json.dumps(line), flush=True)
    capi.shutdown()
