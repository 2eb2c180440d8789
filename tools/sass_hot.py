"""Developer aid: summarise an `ncu --page source --csv --print-source sass` dump.
usage: python tools/sass_hot.py dump.csv [min_exec_fraction]
Prints contiguous regions of SASS grouped by executed-instruction count, with stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]
ci = {h: i for i, h in enumerate(hdr)}
ins = []
for r in rows[start + 1:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    ins.append((r[ci["Source"]].strip(), int(r[ci["Instructions Executed"]] or 0), int(r[ci["# Samples"]] or 0)))
tot_e = sum(e for _, e, _ in ins); tot_s = sum(s for _, _, s in ins)
print(f"{len(ins)} SASS instructions, executed {tot_e/1e6:.1f} M warp-instrs, {tot_s} samples")
# regions: runs with similar execution count (within 25 %)
i = 0
regions = []
while i < len(ins):
    j = i; e0 = ins[i][1]
    while j + 1 < len(ins) and (abs(ins[j + 1][1] - e0) <= 0.25 * max(e0, 1)):
        j += 1
    regions.append((i, j)); i = j + 1
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
for a, b in regions:
    e = sum(x[1] for x in ins[a:b + 1]); s = sum(x[2] for x in ins[a:b + 1])
    if e / tot_e >= thr or s / max(tot_s, 1) >= thr:
        ops = {}
        for src, ee, _ in ins[a:b + 1]:
            op = src.split()[0] if not src.startswith("@") else src.split()[1]
            op = op.split(".")[0]
            ops[op] = ops.get(op, 0) + 1
        top = sorted(ops.items(), key=lambda kv: -kv[1])[:8]
        print(f"[{a:5d}-{b:5d}] n={b-a+1:4d} exec/instr={ins[a][1]/1e6:7.2f}M  exec%={100*e/tot_e:5.1f} samp%={100*s/max(tot_s,1):5.1f}  {top}")
