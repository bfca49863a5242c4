#!/usr/bin/env python
"""Registers, stack and spills per kernel of one translation unit (nvcc -Xptxas -v, cross-compiled for sm_100a; no GPU needed).

    python tools/ptxas_regs.py [kernels_f32.cu] [extra nvcc flags...]
"""
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def main():
    args = sys.argv[1:]
    src = args[0] if args and args[0].endswith(".cu") else "kernels_f32.cu"
    extra = [a for a in args if not a.endswith(".cu")]
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xptxas", "-v", "-c",
           str(ROOT / "paths_b200" / "csrc" / src), "-o", "/dev/null"] + extra
    if src == "kernels_f64.cu":
        cmd.append("-fmad=false")
    out = subprocess.run(cmd, capture_output=True, text=True).stderr
    name = None
    rows = []
    stack = spill = ""
    for line in out.splitlines():
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name).replace("void paths::", "")
            continue
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m:
            stack, spill = m.group(1), f"{m.group(2)}/{m.group(3)}"
            continue
        m = re.search(r"Used (\d+) registers.*?(?:(\d+) bytes smem)?", line)
        if m and name:
            sm = re.search(r"(\d+) bytes smem", line)
            rows.append((name, int(m.group(1)), stack, spill, sm.group(1) if sm else "0"))
            name = None
    print(f"{'kernel':64s} regs stack spill(st/ld) smem")
    for r in rows:
        print(f"{r[0][:64]:64s} {r[1]:4d} {r[2]:>5s} {r[3]:>9s} {r[4]:>6s}")


if __name__ == "__main__":
    main()
