"""Per-opcode table from an ncu report's SASS source page: share of warp instructions, share of stall samples, threads per instruction.

usage: ncu_opcodes.py report.ncu-rep [kernel_substr] [launch_index]
Needs no cubin (unlike ncu_lines.py); shows e.g. how many of a kernel's stall samples sit on ATOM / LDG / BAR.
"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
ksub = sys.argv[2] if len(sys.argv) > 2 else ""
launch = int(sys.argv[3]) if len(sys.argv) > 3 else 0
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
# the source page lists every kernel twice (two views); distinct launches are every other block
b = blocks[launch]
h = b["hdr"]
iS, iI, iT, iSrc = h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("Source")
inst, thr, samp = collections.Counter(), collections.Counter(), collections.Counter()
for r in b["rows"]:
    toks = [t for t in r[iSrc].split() if not t.startswith("@")]
    op = toks[0].split(".")[0] if toks else "?"
    if op in ("DFMA", "DADD", "DMUL", "DSETP", "DMNMX") or ".F64" in r[iSrc]:
        op = "F64"
    if op in ("ATOM", "ATOMG", "RED", "ATOMS"):
        op = "ATOM/RED"
    inst[op] += int(r[iI])
    thr[op] += int(r[iT])
    samp[op] += int(r[iS])
ti, ts = sum(inst.values()), max(sum(samp.values()), 1)
print("%s  launch %d: warp instructions %d, stall samples %d, threads/instruction %.1f" % (b["name"][:90], launch, ti, ts, sum(thr.values()) / max(ti, 1)))
print("%-10s %8s %8s %8s" % ("opcode", "inst %", "samp %", "thr/inst"))
for op, v in sorted(inst.items(), key=lambda kv: -samp[kv[0]])[:18]:
    print("%-10s %8.1f %8.1f %8.1f" % (op, 100.0 * v / ti, 100.0 * samp[op] / ts, thr[op] / max(v, 1)))
