"""Digest of an `ncu --set full` capture of the layer-1 kernel: profiles/<name>_summary.txt and profiles/r02_traffic.json
(the DRAM bytes per launch that bench.py reports as roofline.traffic while the kernel's source is the one profiled).

python tools/make_ncu_digest.py gpurun_out/r02_l1stream.ncu-rep r02_ncu_l1stream
python tools/make_ncu_digest.py gpurun_out/conv_persist.ncu-rep r02_ncu_conv_persist conv     (the C4 conv2d kernel ->
                                                                                              profiles/r02_traffic_conv.json)
"""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes_read.sum.per_second", "dram__bytes_read.sum.pct_of_peak_sustained_elapsed",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum.per_second", "lts__t_sector_hit_rate.pct",
        "launch__registers_per_thread", "launch__block_size", "launch__grid_size", "launch__cluster_size",
        "launch__cluster_max_active", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active"]


def main():
    rep, name = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
    hdr, units, vals = rows[0], rows[1], rows[2]
    m = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    lines = [f"ncu --set full --clock-control none, one launch of: {m['Kernel Name'][0]}", f"report: {os.path.basename(rep)}", ""]
    for k in KEYS:
        if k in m:
            lines.append(f"{k:100s} {m[k][0]:>16s} {m[k][1]}")
    unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}
    rd = float(m["dram__bytes_read.sum"][0].replace(",", "")) * unit[m["dram__bytes_read.sum"][1]]
    wr = float(m["dram__bytes_write.sum"][0].replace(",", "")) * unit[m["dram__bytes_write.sum"][1]]
    conv = len(sys.argv) > 3 and sys.argv[3] == "conv"
    if conv:
        alg = 64 * 112 * 112 * 64 * 2.0 + 64 * 110 * 110 * 64 * 4.0
        lines += ["", f"DRAM bytes per launch: read {rd / 1e6:.1f} MB + write {wr / 1e6:.1f} MB = {(rd + wr) / 1e6:.1f} MB; algorithmic (bf16 NHWC image "
                  f"once + fp32 result once) {alg / 1e6:.1f} MB -> {(rd + wr) / alg:.3f} x (part of the result is still in L2 when the kernel ends)"]
        open(os.path.join(ROOT, "profiles", name + "_summary.txt"), "w").write("\n".join(lines) + "\n")
        src = open(os.path.join(ROOT, "netsdb_b200", "csrc", "nb200_gemm.cu"), "rb").read()
        json.dump({"kernel_source_sha16": hashlib.sha256(src).hexdigest()[:16], "report": os.path.basename(rep),
                   "workload": "conv2d_memory_fusion C4: 64 x 64ch x 112 x 112, 3x3, 64 filters, bf16",
                   "k_conv_persist": {"dram_bytes_read": rd, "dram_bytes_write": wr, "algorithmic_bytes": alg}},
                  open(os.path.join(ROOT, "profiles", "r02_traffic_conv.json"), "w"), indent=1)
        print("\n".join(lines))
        return
    alg = 10000 * 597540 * 8.0 + 1000 * 597540 * 2.0
    lines += ["", f"DRAM bytes per launch: read {rd / 1e9:.2f} GB + write {wr / 1e9:.3f} GB = {(rd + wr) / 1e9:.2f} GB; algorithmic "
              f"(X fp64 once + W1 fp16 once) {alg / 1e9:.2f} GB -> {(rd + wr) / alg:.3f} x",
              "tensor pipe % is the average over all 148 SMs; the 8-CTA clusters occupy 120 of them"]
    open(os.path.join(ROOT, "profiles", name + "_summary.txt"), "w").write("\n".join(lines) + "\n")
    src = open(os.path.join(ROOT, "netsdb_b200", "csrc", "nb200_l1stream.cu"), "rb").read()
    json.dump({"kernel_source_sha16": hashlib.sha256(src).hexdigest()[:16], "report": os.path.basename(rep),
               "workload": "Amazon-14K, 10,000 rows, F16", "k_l1_stream": {"dram_bytes_read": rd, "dram_bytes_write": wr,
                                                                              "algorithmic_bytes": alg}},
              open(os.path.join(ROOT, "profiles", "r02_traffic.json"), "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
