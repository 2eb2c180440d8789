#!/bin/bash
# Developer aid: per-function shares of executed warp instructions.
#   tools/prof_regions.sh <ncu source csv> <nvdisasm -g -c output> [top lines]
python - "$1" "$2" "${3:-25}" <<'PY'
import re, subprocess, sys, glob
src = open(glob.glob('/root/repo/research*/csrc/cg_warp.cuh')[0]).read().split('\n')
funcs = []
for i, l in enumerate(src, 1):
    if re.match(r'^(CG_DEVINL|CG_DEVNOINL|__device__|__global__|inline|warp_kernel)', l):
        name = re.search(r'(\w+)\s*\(', l)
        funcs.append((i, name.group(1) if name else l[:30]))
args = []
for k, (s, n) in enumerate(funcs):
    e = funcs[k + 1][0] - 1 if k + 1 < len(funcs) else len(src)
    args.append(f"{n}:{s}-{e}")
print(subprocess.run(['python', '/root/repo/tools/sass_profile.py', sys.argv[1], sys.argv[2], sys.argv[3]] + args,
                     capture_output=True, text=True).stdout)
PY
