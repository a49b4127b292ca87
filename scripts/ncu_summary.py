#!/usr/bin/env python
"""Prints the key raw metrics and the top stall locations of each kernel in an .ncu-rep file."""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__inst_executed.sum", "smsp__cycles_active.avg"]


def main(rep, topn=14):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("====", r[hdr.index("Kernel Name")][:90])
        for w in WANT:
            if w in hdr:
                print(f"  {w:82s} {r[hdr.index(w)]:>16s} {units[hdr.index(w)]}")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    blocks = [i for i, r in enumerate(rows) if "Source" in r and any("Sampling (All" in c for c in r)]
    for bi, h in enumerate(blocks[:3]):
        end = blocks[bi + 1] if bi + 1 < len(blocks) else len(rows)
        hd = rows[h]
        si = hd.index("Source")
        sa = [j for j, c in enumerate(hd) if "Sampling (All" in c][0]
        data = []
        for k, r in enumerate(rows[h + 1:end]):
            try:
                data.append((int(r[sa]), k, r[si][:100]))
            except (ValueError, IndexError):
                pass
        tot = sum(v for v, _, _ in data) or 1
        print(f"---- kernel instance {bi}: {tot} samples, top {topn} SASS lines (index in listing)")
        for v, k, s in sorted(data, reverse=True)[:topn]:
            print(f"  {100 * v / tot:5.1f}%  #{k:5d}  {s}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 14)
