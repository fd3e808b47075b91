"""Reads an .ncu-rep (ncu -i ... --page raw --csv) and prints one line per captured launch with the metrics the roofline discussion uses."""
import csv
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "dram_rd_MB"), ("dram__bytes_write.sum", "dram_wr_MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"), ("launch__registers_per_thread", "regs"),
        ("smsp__inst_executed_op_shared_atom.sum", "smem_atomics"), ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts"),
        ("smsp__pcsamp_warps_issue_stalled_long_scoreboard", "stall_long_sb"), ("smsp__pcsamp_warps_issue_stalled_short_scoreboard", "stall_short_sb"),
        ("smsp__pcsamp_warps_issue_stalled_lg_throttle", "stall_lg_throttle"), ("smsp__pcsamp_warps_issue_stalled_mio_throttle", "stall_mio"),
        ("smsp__pcsamp_warps_issue_stalled_barrier", "stall_barrier"), ("smsp__pcsamp_sample_count", "samples")]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
col = {}
for key, name in KEYS:
    for i, h in enumerate(hdr):
        if h == key or h.endswith("." + key) or h.endswith(key):
            col[name] = i
            break
ki, gi = hdr.index("Kernel Name"), hdr.index("Grid Size")
print(",".join(["kernel", "grid"] + [n for _, n in KEYS if n in col] + ["units:" + "|".join("%s=%s" % (n, units[col[n]]) for _, n in KEYS if n in col)]))
for r in rows[2:]:
    print(",".join([r[ki].split("(")[0].replace(",", ";"), r[gi].replace(",", "x").replace(" ", "")] + [r[col[n]] for _, n in KEYS if n in col]))
