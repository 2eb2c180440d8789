"""Developer aid: where one stall reason is sampled, per source line.
  python tools/stall_lines.py src.csv disasm.sass stall_no_inst [top]"""
import csv, re, sys
from collections import defaultdict
src_csv, sass, col = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ic, ie = hdr.index(col), hdr.index("Instructions Executed")
vals = [(int(r[ic] or 0), int(r[ie] or 0)) for r in rows[2:] if len(r) > ic]
line, lines = None, []
for l in open(sass):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        line = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", l):
        lines.append(line)
agg = defaultdict(lambda: [0, 0])
for k in range(min(len(lines), len(vals))):
    agg[lines[k]][0] += vals[k][0]
    agg[lines[k]][1] += vals[k][1]
tot = sum(v[0] for v in agg.values())
print(f"{col}: {tot} samples")
for ln, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{v[0]:7d} {100 * v[0] / max(tot, 1):5.1f}%  exec {v[1]:12d}  {ln}")
