"""Text summary of one ncu capture for profiles/: python tools/ncu_summary.py <prefix> <disasm.txt> "<title>"
reads <prefix>_details.txt, <prefix>_raw.csv, <prefix>_sass.csv.gz (exported on the GPU box with ncu -i ... --page ...)."""
import csv, re, subprocess, sys
pre, dis, title = sys.argv[1], sys.argv[2], sys.argv[3]
print(title); print()
keep = re.compile(r"Duration|Registers Per|Executed Ipc|Issue Slots Busy|No Eligible|Theoretical Occ|Achieved Occ|Avg. Active Threads|Avg. Not Pred|DRAM Throughput|Memory Throughput|Block Size|Grid Size|Waves|L1/TEX Hit|L2 Hit|Compute \(SM\)|Issued Warp|Eligible Warps|Active Warps Per|Branch Eff")
for ln in open(pre + "_details.txt"):
    if keep.search(ln): print(ln.rstrip())
rows = list(csv.reader(open(pre + "_raw.csv")))
h, u, v = rows[0], rows[1], rows[2]
keys = {'dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','smsp__inst_executed.sum','smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed','sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','sm__icc_request_hit_rate.pct','gcc__average_cache_request_hit_rate.pct','launch__registers_per_thread'}
print("\nraw counters:")
for n, uu, x in zip(h, u, v):
    if n in keys or ('issue_stalled' in n and 'per_issue_active' in n): print("  %s [%s]: %s" % (n, uu, x))
print("\nexecuted warp instructions by source line (tools/sass_lines.py: ncu SASS page joined with the nvdisasm -g line table):")
sys.stdout.flush()
subprocess.call([sys.executable, "tools/sass_lines.py", pre + "_sass.csv.gz", dis, "30"])
