"""Developer aid: attribute executed instructions / stall samples of an ncu SASS dump to source
lines, using nvdisasm --print-line-info of the same cubin.
usage: python tools/sass_lines.py ncu_sass.csv nvdisasm.sass kernel_substr [top]"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]; ci = {h: i for i, h in enumerate(hdr)}
ins = [(r[ci["Source"]].strip(), int(r[ci["Instructions Executed"]] or 0), int(r[ci["# Samples"]] or 0))
       for r in rows[start + 1:] if len(r) >= len(hdr) and r[0].startswith("0x")]
# nvdisasm: walk the text section of the kernel and the device functions after it, in order
lines = open(sys.argv[2]).read().split("\n")
key = sys.argv[3]
sec = [i for i, l in enumerate(lines) if l.startswith("//--------------------- .text.") and key in l]
assert sec, "kernel not found"
cur = None; seq = []
for l in lines[sec[0]:]:
    if l.startswith("//--------------------- ") and ".text." not in l and seq: break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        inl = re.search(r'inlined at "([^"]+)", line (\d+)', m.group(3))
        cur = (m.group(1).split("/")[-1], int(m.group(2)), (inl.group(1).split("/")[-1], int(inl.group(2))) if inl else None)
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m: seq.append((cur, m.group(2).strip()))
print(len(ins), "ncu instrs;", len(seq), "nvdisasm instrs")
n = min(len(ins), len(seq))
# sanity: opcode agreement
agree = sum(1 for i in range(n) if ins[i][0].split()[0].split(".")[0] == seq[i][1].split()[0].split(".")[0] or ins[i][0].startswith("@") == seq[i][1].startswith("@"))
print("opcode agreement", agree / n)
agg = {}
for i in range(n):
    k = seq[i][0][:2] if seq[i][0] else ("?", 0)
    a = agg.setdefault(k, [0, 0, 0]); a[0] += ins[i][1]; a[1] += ins[i][2]; a[2] += 1
tot = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
src_cache = {}
def text(f, ln):
    import glob
    if f not in src_cache:
        c = glob.glob("/root/repo/**/" + f, recursive=True)
        src_cache[f] = open(c[0]).read().split("\n") if c else []
    s = src_cache[f]
    return s[ln - 1].strip()[:90] if 0 < ln <= len(s) else ""
top = int(sys.argv[4]) if len(sys.argv) > 4 else 60
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0]:18s}:{k[1]:4d} {100*a[0]/tot:5.1f}% exec {100*a[1]/max(ts,1):5.1f}% samp n={a[2]:4d}  {text(*k)}")
# optional grouping by line ranges of one file: env CG_GROUPS="name:lo-hi,name:lo-hi" (file = first arg's main .cuh)
import os
if os.environ.get("CG_GROUPS"):
    fname = os.environ.get("CG_GROUP_FILE", "cg_pairstream.cuh")
    groups = []
    for g in os.environ["CG_GROUPS"].split(","):
        nm, rg = g.split(":"); lo, hi = rg.split("-"); groups.append((nm, int(lo), int(hi)))
    acc = {}
    for k, a in agg.items():
        nm = k[0]
        if k[0] == fname:
            nm = next((g[0] for g in groups if g[1] <= k[1] <= g[2]), "other:" + fname)
        b = acc.setdefault(nm, [0, 0]); b[0] += a[0]; b[1] += a[1]
    print("---- groups")
    for nm, b in sorted(acc.items(), key=lambda kv: -kv[1][1]):
        print(f"{nm:28s} {100*b[0]/tot:5.1f}% exec {100*b[1]/max(ts,1):5.1f}% samp")
