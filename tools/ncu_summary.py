"""Summarise an ncu raw CSV page: one line per kernel with the metrics B200_PROFILING.md names."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"),
        ("dram__bytes_write.sum", "dram_wr"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"), ("launch__registers_per_thread", "regs"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conf"),
        ("smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "st_long_sb"),
        ("smsp__warp_issue_stalled_barrier_per_warp_active.pct", "st_barrier")]
idx = [(hdr.index(w), n) for w, n in want if w in hdr]
print(" | ".join(n for _, n in idx))
for r in rows[2:]:
    out = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.replace("b200::", "").replace("void ", "")[:44].ljust(44)
        else:
            try:
                v = f"{float(v.replace(',', '')):.4g}{units[i][:5]}"
            except ValueError:
                pass
        out.append(v)
    print(" | ".join(out))
