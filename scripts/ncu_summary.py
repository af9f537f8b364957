"""Summarise an .ncu-rep (read on the CPU box): python scripts/ncu_summary.py file.ncu-rep [regex...]"""
import csv
import io
import re
import subprocess
import sys

DEFAULT = [r"gpu__time_duration\.sum", r"dram__bytes_(read|write)\.sum$", r"dram__throughput\.avg\.pct",
           r"gpu__dram_throughput", r"dram__cycles_active", r"sm__warps_active\.avg\.pct",
           r"launch__registers_per_thread", r"launch__shared_mem_per_block_dynamic", r"launch__grid_size",
           r"sm__throughput\.avg\.pct", r"lts__throughput\.avg\.pct", r"l1tex__throughput\.avg\.pct",
           r"sm__pipe_tensor.*cycles_active", r"sm__inst_executed_pipe_tensor", r"smsp__inst_executed\.sum$",
           r"lts__t_sector_hit_rate", r"lts__t_bytes\.sum$", r"sm__cycles_elapsed\.max",
           r"smsp__average_warp.*issue_stalled.*ratio$|smsp__average_warps_issue_stalled.*per_issue_active"]


def main():
    rep = sys.argv[1]
    pats = [re.compile(p) for p in (sys.argv[2:] or DEFAULT)]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("==", r[hdr.index("Kernel Name")][:90], "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
        for i, h in enumerate(hdr):
            if any(p.search(h) for p in pats):
                print(f"   {h} [{units[i]}] = {r[i]}")


if __name__ == "__main__":
    main()
