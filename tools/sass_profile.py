"""Developer aid: join an ncu SASS source page (ncu -i X.ncu-rep --page source --csv) with nvdisasm -g line
info and print executed warp instructions per source line / per function-of-origin.
  python tools/sass_profile.py src.csv disasm.sass [top]"""
import csv
import re
import sys
from collections import defaultdict

src_csv, sass, top = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ia, ii, it, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
execd = [(int(r[ii] or 0), int(r[it] or 0), int(r[isamp] or 0), r[1]) for r in rows[2:] if len(r) > ii]
# line info per instruction, in order
line = None
inl = None
lines = []
for l in open(sass):
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        line = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", l):
        lines.append(line)
n = min(len(lines), len(execd))
per_line = defaultdict(lambda: [0, 0, 0, 0])
total = 0
for k in range(n):
    e = execd[k]
    per_line[lines[k]][0] += e[0]
    per_line[lines[k]][1] += e[1]
    per_line[lines[k]][2] += e[2]
    per_line[lines[k]][3] += 1
    total += e[0]
print(f"instructions: static {n}, executed {total:.3e}")
print("   exec        share  avg_thr  samples  static   file:line")
for ln, v in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{v[0]:12d} {100 * v[0] / total:6.2f}%  {v[1] / max(v[0], 1):5.1f}  {v[2]:7d}  {v[3]:5d}   {ln}")

# per function of origin (line ranges of csrc/cg_warp.cuh given as name:first-last on the command line after `top`)
ranges = []
for arg in sys.argv[4:]:
    name, r = arg.split(":")
    a, b = r.split("-")
    ranges.append((name, int(a), int(b)))
if ranges:
    agg = defaultdict(lambda: [0, 0, 0])
    for ln, v in per_line.items():
        key = "other:" + (ln[0] if ln else "?")
        if ln and ln[0] == "cg_warp.cuh":
            for name, a, b in ranges:
                if a <= ln[1] <= b:
                    key = name
                    break
        agg[key][0] += v[0]; agg[key][1] += v[2]; agg[key][2] += v[3]
    print("\nper region:      exec    share  samples static")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{k:28s} {v[0]:12d} {100 * v[0] / total:6.2f}% {v[1]:7d} {v[2]:6d}")
