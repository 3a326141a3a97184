#!/usr/bin/env python
"""Per-source-line summary of an ncu report (needs -lineinfo + --import-source on):
   python tools/ncu_lines.py report.ncu-rep [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
SEL = ["--launch-skip", skip, "--launch-count", "1"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"] + SEL, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, val = rows[0], rows[2]
want = ["gpu__time_duration.sum", "smsp__inst_executed.sum ", "smsp__issue_active.avg.pct", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "lts__t_bytes.sum ",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread ", "sm__warps_active.avg.pct",
        "smsp__average_warps_issue_stalled", "lts__t_sectors_srcunit_tex_op_read.sum ", "sm__throughput.avg.pct", "l1tex__throughput.avg.pct",
        "lts__throughput.avg.pct", "smsp__inst_executed_op_shared", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum "]
for h, u, v in zip(hdr, rows[1], val):
    if any((h + " ").startswith(w) or w.strip() == h for w in want):
        print("%-90s %-12s %s" % (h, u, v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"] + SEL, capture_output=True, text=True).stdout
out, fname = {}, None
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if len(r) > 10 and r[0].isdigit() and r[2] == "-":
        try:
            inst, thr, samp = int(r[7]), int(r[8]), int(r[6])
        except ValueError:
            continue
        o = out.setdefault((fname, int(r[0])), [0, 0, 0, r[1].strip()[:100]])
        o[0] += inst; o[1] += thr; o[2] += samp
tot = sum(v[0] for v in out.values()) or 1
ts = sum(v[2] for v in out.values()) or 1
print("total warp instructions %d, stall samples %d" % (tot, ts))
for k, v in sorted(out.items(), key=lambda kv: -kv[1][2])[:top]:
    print("%5.1f%%inst %5.1f%%samp thr/inst %4.1f  %s:%d  %s" % (100.0 * v[0] / tot, 100.0 * v[2] / ts, v[1] / max(v[0], 1), k[0][:14], k[1], v[3]))
