#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics + executed-instruction histogram."""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_bytes.sum", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "lts__t_sectors_srcunit_tex_op_write.sum", "lts__t_sectors_op_write.sum", "lts__t_sectors_op_read.sum"]
for i, h in enumerate(hdr):
    if h in want or h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
        vals = [r[i] for r in data]
        print(f"{h} [{units[i]}]: {vals}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
k, hdr, data = 0, None, []
for r in rows:
    if r and r[0] == "Kernel Name":
        k += 1
        continue
    if r and r[0] == "Address":
        hdr = r
        continue
    if k == 1 and hdr and len(r) == len(hdr):
        data.append(r)
if data:
    ie, isrc = hdr.index("Instructions Executed"), hdr.index("Source")
    tot = sum(int(r[ie]) for r in data)
    ops = collections.Counter()
    for r in data:
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[isrc])
        ops[m.group(2).split(".")[0] if m else "?"] += int(r[ie])
    print("executed warp instructions (launch 1):", tot)
    print("  ".join(f"{op}:{100 * c / tot:.1f}%" for op, c in ops.most_common(22)))
