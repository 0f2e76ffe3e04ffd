#!/usr/bin/env python
"""Summarise an .ncu-rep (run here, no GPU needed): headline counters + hottest source lines.
usage: python scripts/ncu_summary.py gpurun_out/x.ncu-rep [top_n]
       python scripts/ncu_summary.py gpurun_out/x.ncu-rep --emit-constants HASHES.json CONFIGS FRAMES SUMMARY.txt > profiles/ksweep_constants.json
--emit-constants writes the per-launch counters bench.py's roofline uses (executed FP32 FLOP, warp instructions, DRAM bytes of
ONE k_sweep launch) together with the SHA-256 of the kernel sources the capture was taken from (HASHES.json = the output of
`python bench.py --print-source-hash` ON THE GPU BOX in the same gpurun call as the capture); bench.py refuses the constants
when the working tree's sources hash differently."""
import csv, json, subprocess, sys, io

rep = sys.argv[1]
if len(sys.argv) > 2 and sys.argv[2] == "--emit-constants":
    hashes = json.load(open(sys.argv[3]))
    configs, frames, summary = int(sys.argv[4]), int(sys.argv[5]), sys.argv[6]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2]
    def val(k, scale={"Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Gbyte": 1e9, "ms": 1.0, "us": 1e-3, "inst": 1.0, "": 1.0}):
        i = hdr.index(k)
        return float(data[i].replace(",", "")) * scale.get(units[i], 1.0)
    out = {"kernel": data[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "k_sweep", "configs_per_launch": configs, "frames": frames,
           "warp_inst": val("smsp__inst_executed.sum"),
           "flop": val("smsp__sass_thread_inst_executed_op_fadd_pred_on.sum") + val("smsp__sass_thread_inst_executed_op_fmul_pred_on.sum")
                   + 2 * val("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum"),
           "dram_read_bytes": val("dram__bytes_read.sum"), "dram_write_bytes": val("dram__bytes_write.sum"),
           "gpu_time_ms_under_ncu": val("gpu__time_duration.sum"), "source_sha256": hashes, "ncu_report": rep, "summary": summary}
    print(json.dumps(out, indent=1))
    sys.exit(0)
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum",
        "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
for k in keys:
    if k in hdr:
        i = hdr.index(k)
        print(f"{k} [{units[i]}] = {[r[i] for r in data]}")
for h in hdr:
    if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
        i = hdr.index(h)
        v = float(data[0][i])
        if v > 0.15:
            print(f"  stall {h.split('stalled_')[1].split('_per')[0]}: {v:.2f}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-count", "1"], capture_output=True, text=True).stdout
cur, hd, agg = None, None, []
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 5 and r[0] == "Line No":
        hd = r
    elif hd and len(r) == len(hd) and r[0] != "":
        agg.append((cur, int(r[0]), r[1].strip()[:80], int(r[6]), int(r[7]), int(r[8])))
ts, ti = sum(a[3] for a in agg) or 1, sum(a[4] for a in agg) or 1
print(f"source: samples {ts} warp-inst {ti}")
for a in sorted(agg, key=lambda a: -a[3])[:top]:
    print(f"{a[0]:20s}{a[1]:5d} samp {100*a[3]/ts:5.1f}% inst {100*a[4]/ti:5.1f}% thr/inst {a[5]/max(1,a[4]):5.1f} | {a[2]}")
