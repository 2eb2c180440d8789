"""In-tree build of libcanarygate.so (hand-written sm_100a CUDA behind a C ABI).

nvcc cross-compiles here without a GPU; the .so is git-ignored but travels to the
GPU box with the gpurun snapshot.  -fmad=false: the reference computes in Python
floats, one IEEE rounding per operation, so no multiply-add may be contracted.
The streaming kernel is instantiated per (element type, quantile count) in separate
translation units that compile in parallel.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libcanarygate.so")
HEADERS = ["cg_math.cuh", "cg_ptx.cuh", "cg_general.cuh", "cg_rows.cuh", "cg_warp.cuh", "cg_irregular.cuh", "cg_irows.cuh", "cg_synth.cuh", "cg_store.cuh",
           os.path.join(ROOT, "include", "canary_gate.h")]

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ARCH + ["-O3", "-lineinfo", "-std=c++17", "-fmad=false", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _units():
    # CG_ABLATE=1: developer build with the tuning knobs (ring depth, CTAs per SM) read from the environment
    extra = ["-DCG_ABLATE"] if os.environ.get("CG_ABLATE") else []
    if os.environ.get("CG_MIN_CTAS"):                # developer build: register budget of the streaming kernel
        extra += [f"-DCG_MIN_CTAS={int(os.environ['CG_MIN_CTAS'])}"]
    if os.environ.get("CG_ROWS_PER_USE"):            # developer build: rows per ring slot
        extra += [f"-DCG_ROWS_PER_USE={int(os.environ['CG_ROWS_PER_USE'])}"]
    if os.environ.get("CG_MAXNREG"):                 # developer build: explicit register cap instead of a CTA count
        extra += [f"-DCG_MAXNREG={int(os.environ['CG_MAXNREG'])}"]
    if os.environ.get("CG_DEFS"):                    # developer build: experimental switches, comma separated
        extra += [f"-D{d}" for d in os.environ["CG_DEFS"].split(",") if d]
    if os.environ.get("CG_G"):                       # developer build: periods of the tick grid per lane (1, 2 or 4)
        extra += [f"-DCG_G={int(os.environ['CG_G'])}"]
    units = [("cg_api", "cg_api.cu", extra)]
    for tname, tchar in (("double", "d"), ("float", "f")):
        for nq in (1, 2, 3, 4):
            name = f"cg_launch_warp_{tchar}{nq}"
            defs = [f"-DCG_INST_T={tname}", f"-DCG_INST_NQ={nq}", f"-DCG_INST_NAME={name}"] + extra
            if os.environ.get("CG_CHECKED"):                 # developer build: our own bounds checks (tools/run_checked.py)
                defs += ["-DCG_CHECKED"] + (["-DCG_CHECK_EXPORT"] if name == "cg_launch_warp_d1" else [])
            units.append((name, "cg_warp_inst.cu", defs))
    return units


def _nvcc() -> str:
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: libcanarygate.so cannot be built")


def _newest_dep(src: str) -> float:
    deps = [os.path.join(CSRC, src)] + [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in HEADERS]
    deps.append(os.path.abspath(__file__))
    return max(os.path.getmtime(d) for d in deps if os.path.exists(d))


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(_newest_dep(src) > t for _, src, _ in _units())


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    nvcc = _nvcc()
    os.makedirs(OBJ, exist_ok=True)
    log_path = os.path.join(HERE, "build.log")
    logs = []

    def compile_one(unit):
        name, src, defs = unit
        obj = os.path.join(OBJ, name + ".o")
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= _newest_dep(src):
            return obj, 0, ""
        cmd = [nvcc] + CFLAGS + defs + ["-c", os.path.join(CSRC, src), "-o", obj]
        res = subprocess.run(cmd, capture_output=True, text=True)
        return obj, res.returncode, " ".join(cmd) + "\n" + res.stdout + res.stderr

    with ThreadPoolExecutor(max_workers=min(9, os.cpu_count() or 1)) as ex:
        results = list(ex.map(compile_one, _units()))
    objs = []
    for obj, rc, log in results:
        logs.append(log)
        if rc != 0:
            with open(log_path, "w") as f:
                f.write("\n".join(logs))
            raise RuntimeError("nvcc failed:\n" + log)
        objs.append(obj)
    cmd = [nvcc] + ARCH + ["--shared", "-cudart", "shared", "-o", LIB] + objs
    res = subprocess.run(cmd, capture_output=True, text=True)
    logs.append(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    with open(log_path, "w") as f:
        f.write("\n".join(logs))
    if verbose:
        print("\n".join(logs))
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose=True))
