"""Print selected metrics of every kernel in an .ncu-rep (read with `ncu -i ... --page raw --csv`)."""
import csv, subprocess, sys
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__grid_size"]
rep = sys.argv[1]
extra = sys.argv[2:]
if rep.endswith(".csv"):      # the raw page saved as csv on the GPU box
    rows = list(csv.reader([l for l in open(rep) if not l.startswith("==")]))
else:
    rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
h = rows[0]
for v in rows[2:]:
    print("==", v[h.index("Kernel Name")][:80])
    for i, k in enumerate(h):
        if k in KEYS or "issue_stalled" in k and k.endswith("per_issue_active.ratio") or any(e in k for e in extra):
            print(f"  {k:90s} {rows[1][i]:12s} {v[i]}")
