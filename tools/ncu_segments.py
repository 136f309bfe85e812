"""Summarise an .ncu-rep of one kernel: headline metrics + instruction/stall-sample counts per barrier-delimited segment."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
keep = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum', 'launch__grid_size',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sectors_srcunit_tex_op_read.sum']
grid = 1
for h, u, v in zip(hdr, units, vals):
    if h in keep or ('issue_stalled' in h and 'per_issue_active' in h and float(v.replace(',', '') or 0) > 0.3):
        print(f"{h},{u},{v}")
    if h == 'launch__grid_size': grid = int(float(v.replace(',', '')))
    if h == 'launch__block_size': blk = int(float(v.replace(',', '')))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
ia, isrc, ist = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
data = rows[2:]
tot = sum(int(r[ia]) for r in data); tots = sum(int(r[ist]) for r in data)
wpb = blk // 32
print(f"# total warp-inst {tot}, per warp per CTA {tot / grid / wpb:.1f}, samples {tots}")
seg = acc = accs = start = 0
for i, r in enumerate(data):
    acc += int(r[ia]); accs += int(r[ist])
    if 'BAR.SYNC' in r[isrc] or i == len(data) - 1:
        print(f"seg {seg}: sass[{start}:{i}] inst/warp={acc / grid / wpb:.1f} ({acc / tot:.3f}) samples={accs / tots:.3f}")
        seg += 1; acc = accs = 0; start = i + 1
