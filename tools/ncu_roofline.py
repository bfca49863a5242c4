"""ncu per-launch metric CSV (`--csv --log-file`, one row per launch and metric) -> profiles/traffic.json and profiles/issue.json.

    python tools/ncu_roofline.py SCENE capture.csv probe.json GIT_HEAD

Per kernel class (camera, extend, shade, shadow, accumulate): launches, summed gpu__time_duration, DRAM bytes read + written, and the
time-weighted issue-slot utilisation and threads-per-instruction.  The files carry the git head the capture was started from and a hash
of the kernel sources, so that bench.py can refuse numbers that belong to other code."""
import csv
import hashlib
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def source_hash():
    h = hashlib.sha256()
    for p in sorted((ROOT / "paths_b200" / "csrc").glob("*")):
        if p.suffix in (".cu", ".cuh", ".h", ".cpp"):
            h.update(p.name.encode())
            h.update(p.read_bytes())
    return h.hexdigest()[:16]


def classify(name):
    for key, cls in (("k_camera_extend", "camera"), ("k_extend", "extend"), ("k_shade", "shade"), ("k_shadow", "shadow"), ("k_accumulate", "accumulate"),
                     ("k_raygen", "raygen"), ("k_tail", "tail")):
        if key in name:
            return cls
    return "other"


def main():
    scene, csv_path, probe_path, head = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4]
    rows = {}
    with open(csv_path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for rec in csv.DictReader(lines):
        i = int(rec["ID"])
        e = rows.setdefault(i, {"name": rec["Kernel Name"]})
        try:
            e[rec["Metric Name"]] = float(rec["Metric Value"].replace(",", ""))
        except ValueError:
            pass
    probe = json.loads(Path(probe_path).read_text().strip().splitlines()[-1])
    per = {}
    for i in sorted(rows):
        e = rows[i]
        c = per.setdefault(classify(e["name"]), {"launches": 0, "ns": 0.0, "dram_bytes": 0.0, "issue_w": 0.0, "lanes_w": 0.0, "kernels": set()})
        ns = e.get("gpu__time_duration.sum", 0.0)
        c["launches"] += 1
        c["ns"] += ns
        c["dram_bytes"] += e.get("dram__bytes_read.sum", 0.0) + e.get("dram__bytes_write.sum", 0.0)
        c["issue_w"] += ns * e.get("smsp__issue_active.avg.pct_of_peak_sustained_active", 0.0)
        c["lanes_w"] += ns * e.get("smsp__thread_inst_executed_per_inst_executed.ratio", 0.0)
        c["kernels"].add(e["name"].split("(")[0].replace("void ", ""))
    total_ns = sum(c["ns"] for c in per.values()) or 1.0
    out_classes = {}
    for k, c in per.items():
        ns = c["ns"] or 1.0
        out_classes[k] = {"launches": c["launches"], "ms": c["ns"] / 1e6, "share_of_pass": c["ns"] / total_ns,
                          "dram_bytes": c["dram_bytes"], "dram_bytes_per_launch": c["dram_bytes"] / max(c["launches"], 1),
                          "dram_gbs": c["dram_bytes"] / ns, "issue_active_pct": c["issue_w"] / ns, "lanes_per_instruction": c["lanes_w"] / ns,
                          "issue_x_lanes": c["issue_w"] / ns / 100.0 * c["lanes_w"] / ns / 32.0, "kernels": sorted(c["kernels"])}
    stamp = {"git_head": head, "source_hash": source_hash(), "scene": scene, "spp_per_pass": 16,
             "how": "tools/capture_roofline.sh: ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,"
                    "smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none over one "
                    "16-spp 1080p pass of tools/roofline_probe.py (launch times under ncu are cold-cache and serialised: shares, not absolutes)",
             "probe": probe}
    for fname in ("traffic.json", "issue.json"):
        p = ROOT / "profiles" / fname
        try:
            cur = json.loads(p.read_text())
        except (OSError, ValueError):
            cur = {}
        cur = {k: v for k, v in cur.items() if isinstance(v, dict) and "classes" in v}  # drop entries of the old format
        cur[scene] = dict(stamp, classes=out_classes)
        p.write_text(json.dumps(cur, indent=1) + "\n")
    dom = max((k for k in out_classes if k != "other"), key=lambda k: out_classes[k]["ms"])
    print(json.dumps({"scene": scene, "dominant": dom, **{k: {m: round(v, 4) if isinstance(v, float) else v for m, v in c.items() if m != "kernels"}
                                                            for k, c in out_classes.items()}}))


if __name__ == "__main__":
    main()
