#!/usr/bin/env python
"""profiles/roofline_inputs.json from an `ncu --set full` capture of ONE encoder-form backward call (the
dense-window kernel and the per-row kernel in split mode), keyed by a hash of kinet_b200/csrc so that bench.py can
tell when the capture no longer describes the kernels it measures.

    python tools/roofline_inputs.py gpurun_out/prof_enc_bwd.ncu-rep profiles/r2_ncu_full_C2enc_bf16_backward.txt
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import csrc_hash  # noqa: E402

rep, summary = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]


def col(r, name):
    return float(r[hdr.index(name)].replace(",", "")) if name in hdr and r[hdr.index(name)] not in ("", "n/a") else None


dense = [r for r in rows[2:] if "dense_kernel" in r[hdr.index("Kernel Name")]]
prow = [r for r in rows[2:] if "msda_backward_kernel" in r[hdr.index("Kernel Name")]]
assert dense and prow, "capture must contain the dense-window kernel and the per-row kernel"
d, p = dense[0], prow[0]


def to_bytes(r, name):
    v, u = col(r, name), units[hdr.index(name)]
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]


def to_us(r):
    v, u = col(r, "gpu__time_duration.sum"), units[hdr.index("gpu__time_duration.sum")]
    return v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}[u]


traffic = sum(to_bytes(r, "dram__bytes_read.sum") + to_bytes(r, "dram__bytes_write.sum") for r in (d, p))
res = {
    "csrc_hash": csrc_hash(),
    "source": os.path.relpath(summary, ROOT),
    "workload": "C2_enc bf16 value / f32 loc+attn, one backward call (structured sampling locations)",
    "traffic": traffic,
    "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum of the dense-window kernel + the per-row kernel",
    "red_sectors": sum(col(r, "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum") for r in (d, p)),
    "dense_us": to_us(d), "rows_us": to_us(p),
    "l1_data_pipe_pct": col(d, "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
    "tensor_pipe_pct_dense": col(d, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    "issue_active_pct_dense": col(d, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "dram_pct_dense": col(d, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    "dram_pct_rows": col(p, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    "lts_pct_rows": col(p, "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
}
with open(os.path.join(ROOT, "profiles", "roofline_inputs.json"), "w") as f:
    json.dump(res, f, indent=1)
    f.write("\n")
print(json.dumps(res, indent=1))
# human-readable summary next to it
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum", "l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
        "sm__cycles_active.avg", "sm__cycles_elapsed.max", "smsp__sass_inst_executed_op_tmem_ldt.sum"]
with open(summary, "w") as f:
    f.write("ncu --set full --clock-control none, one encoder-form backward call, C2_enc bf16 (N=2, Lq=S=22223, M=8, D=32, L=4, P=4)\n")
    f.write("kernel sources: csrc hash %s\n\n" % res["csrc_hash"])
    for r in (d, p):
        f.write("== %s\n" % r[hdr.index("Kernel Name")][:140])
        for w in WANT:
            if w in hdr:
                f.write("%-88s %18s %s\n" % (w, r[hdr.index(w)], units[hdr.index(w)]))
        f.write("\n")
print("wrote", summary)
