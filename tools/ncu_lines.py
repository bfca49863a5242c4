"""Joins an ncu SASS source page with nvdisasm line info: per source line, stall samples / instructions / SIMT efficiency.

usage: ncu_lines.py report.ncu-rep kernel_substr [launch_index] [cubin]
"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

rep, ksub = sys.argv[1], sys.argv[2]
launch = int(sys.argv[3]) if len(sys.argv) > 3 else 0
cubin = sys.argv[4] if len(sys.argv) > 4 else "/tmp/cub/kernels_f32.sm_100a.cubin"

txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(txt)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and row and row[0] == "Address":
        cur["hdr"] = row
    elif cur is not None and cur["hdr"] and len(row) == len(cur["hdr"]):
        cur["rows"].append(row)
blocks = [b for b in blocks if ksub in b["name"]]
b = blocks[launch]
h = b["hdr"]
iS, iI, iT, iSrc = h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("Source")

dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
# find the function section for the kernel
lines = dis.splitlines()
rows = b["rows"]
key = ksub.split("<")[0].split("(")[0].split("::")[-1]
starts = [(i, re.match(r"\s*\.text\.(\S+):", ln).group(1)) for i, ln in enumerate(lines) if re.match(r"\s*\.text\.(\S+):", ln)]
starts = [(i, m) for i, m in starts if key in m]
if not starts:
    sys.exit("kernel not found in cubin")


def disassemble(start):
    seq, curline = [], ("?", 0)
    for ln in lines[start + 1:]:
        if re.match(r"\s*\.text\.", ln) or ln.strip().startswith(".section"):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            curline = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            seq.append((curline, m.group(2)))
    return seq


# several template instantiations share the name: take the one whose instruction count matches the profiled kernel
cands = [(abs(len(disassemble(i)) - len(rows)), i, m) for i, m in starts]
_, start, mangled = min(cands)
seq = disassemble(start)
if len(seq) != len(rows):
    print(f"warning: {len(seq)} disassembled instructions vs {len(rows)} profiled", file=sys.stderr)
agg = defaultdict(lambda: [0, 0, 0])
for (loc, _), r in zip(seq, rows):
    a = agg[loc]
    a[0] += int(r[iS] or 0)
    a[1] += int(r[iI] or 0)
    a[2] += int(r[iT] or 0)
totS = sum(a[0] for a in agg.values()) or 1
totI = sum(a[1] for a in agg.values()) or 1
print(f"kernel {b['name'][:80]}  launch {launch}: samples {totS}, warp instructions {totI}, SIMT eff {sum(a[2] for a in agg.values()) / totI:.1f}/32")
src_cache = {}
for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    f, n = loc
    if f not in src_cache:
        try:
            src_cache[f] = open("/root/repo/paths_b200/csrc/" + f).read().splitlines()
        except OSError:
            src_cache[f] = []
    text = src_cache[f][n - 1].strip()[:90] if 0 < n <= len(src_cache[f]) else ""
    print(f"{f}:{n:<4d} inst {a[1] / totI:6.1%} samp {a[0] / totS:6.1%} eff {a[2] / max(a[1], 1):5.1f}  {text}")
