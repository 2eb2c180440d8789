"""Developer aid: print the SASS of an address range with executed counts per pair from an ncu source csv.
  python tools/sass_annot.py src.csv disasm.sass pairs 0xba00 0xdd00"""
import csv, re, sys
src_csv, sass, pairs, lo, hi = sys.argv[1], sys.argv[2], float(sys.argv[3]), int(sys.argv[4], 16), int(sys.argv[5], 16)
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ii, it, isamp = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
ex = [(int(r[ii] or 0), int(r[it] or 0), int(r[isamp] or 0)) for r in rows[2:] if len(r) > ii]
k = 0
line = ""
for l in open(sass):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        line = m.group(1).split("/")[-1] + ":" + m.group(2)
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        addr = int(m.group(1), 16)
        if lo <= addr < hi and k < len(ex):
            e = ex[k]
            print(f"{addr:06x} {e[0] / pairs:8.1f} {e[1] / max(e[0], 1):5.1f} {e[2]:5d}  {m.group(2)[:70]:70s} {line}")
        k += 1
