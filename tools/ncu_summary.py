"""Developer aid: key counters of one ncu capture.  python tools/ncu_summary.py X.ncu-rep [pairs]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
pairs = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem',
        'launch__occupancy_limit_registers', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'smsp__average_warp_latency_per_inst_issued.ratio', 'smsp__inst_executed_op_local_ld.sum',
        'smsp__inst_executed_op_local_st.sum', 'launch__grid_size', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'launch__shared_mem_per_block_dynamic']
d = dict(zip(hdr, zip(units, vals)))
for k in want:
    if k in d:
        print(f"{k:70s} {d[k][1]} {d[k][0]}")
for k in hdr:
    if 'average_warps_issue_stalled' in k and k.endswith('per_issue_active.ratio') and float(d[k][1] or 0) >= 0.05:
        print(f"  stall {k.split('issue_stalled_')[1].split('_per_issue')[0]:28s} {float(d[k][1]):.3f}")
if pairs:
    print(f"warp instructions per pair: {float(d['smsp__inst_executed.sum'][1].replace(',', '')) / pairs:.0f}")
