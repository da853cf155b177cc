#!/usr/bin/env python
"""Text report of one .ncu-rep (made on the GPU box, where the report is; only the text comes
back): per launch the roofline metrics, the warp-stall breakdown from the PC samples and the
source lines with the most samples.

Usage: python profiles/ncu_report.py x.ncu-rep [top_lines]"""
import csv
import io
import re
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores"]


def run(args):
    return subprocess.run(["ncu", "-i"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    rows = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for li, r in enumerate(rows[2:]):
        print("=" * 100)
        print("launch %d kernel: %s" % (li, r[idx["Kernel Name"]].split("(")[0]))
        for w in WANT:
            if w in idx:
                print("  %-62s %-16s %s" % (w, units[idx[w]], r[idx[w]]))
        st = {}
        for h, v in zip(hdr, r):
            m = re.match(r"smsp__pcsamp_warps_issue_stalled_(\w+)$", h)
            if m and not h.endswith("not_issued") and v not in ("", "n/a"):
                st[m.group(1)] = float(v)
        tot = sum(st.values()) or 1.0
        print("  stalls: " + ", ".join("%s %.0f%%" % (k, 100 * v / tot)
                                       for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:7]))
        # hottest source lines of this launch
        src = run([rep, "--page", "source", "--csv", "--print-source", "cuda",
                   "--launch-skip", str(li), "--launch-count", "1"])
        srows = list(csv.reader(io.StringIO(src)))
        h = None
        for k, row in enumerate(srows):
            if "Source" in row and any("Samp" in c for c in row):
                h = k
                break
        if h is None:
            print("  (no source page)")
            continue
        sh = srows[h]
        ci = {c: i for i, c in enumerate(sh)}
        samp = next((c for c in sh if c.startswith("# Samples") or c.startswith("Warp Stall Sampling (All")), None)
        inst = next((c for c in sh if c == "Instructions Executed"), None)
        if samp is None:
            print("  (no sample column: %s)" % sh[:12])
            continue
        body = []
        for row in srows[h + 1:]:
            if len(row) != len(sh):
                continue
            try:
                s = float(row[ci[samp]] or 0)
            except ValueError:
                continue
            body.append((s, row))
        total = sum(s for s, _ in body) or 1.0
        body.sort(key=lambda t: -t[0])
        print("  hottest lines (%s, total %.0f):" % (samp, total))
        for s, row in body[:top]:
            ie = row[ci[inst]] if inst else ""
            line = row[ci["#"]] if "#" in ci else ""
            print("   %5.1f%%  inst %-10s L%-5s %s" % (100 * s / total, ie, line, row[ci["Source"]].strip()[:110]))


if __name__ == "__main__":
    main()
