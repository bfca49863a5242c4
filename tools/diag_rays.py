"""Development probe (GPU box): prints what the oracle knows about the fixed rays whose primitive id differs between a production kernel and
the oracle without being a tie or an edge graze.   python tools/diag_rays.py environment [kernel]"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT / "tools"))
import make_assets  # noqa: E402
import oracle  # noqa: E402
import paths_b200 as pb  # noqa: E402
import test_gpu_production_rays as T  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "environment"
kernel = int(sys.argv[2]) if len(sys.argv) > 2 else 32
make_assets.ensure("dragon", "hdri", "table")
case = T.Case(oracle, name)
gsc = pb.Scene.load(f"scenes/{name}.yml", width=T.W, height=T.H)
for label, rays in (("camera", case.camera), ("incoherent", case.incoherent)):
    o, d, p0, t0, b0 = rays
    p1, t1 = gsc.find_intersection(o, d, precision=kernel)
    p64, t64 = gsc.find_intersection(o, d, precision=64)
    bad = T.explain_id_differences(oracle, case.osc, gsc, o, d, p0, b0, p1)
    print(label, "ids differ:", int((p0 != p1).sum()), "unexplained:", len(bad), "f64 kernel differs from oracle:", int((p0 != p64).sum()))
    for i in bad[:12]:
        print(f" ray {i}: o {o[i]} d {d[i]}")
        for tag, p in (("oracle", p0[i]), ("gpu", p1[i])):
            if p == oracle.MISS:
                print(f"   {tag}: MISS")
                continue
            is_tri, geom = gsc.prim_geometry(p)
            tt = case.osc.intersect_prim(p, o[i], d[i])
            if is_tri:
                tb, mb = T.tri_bary(geom, o[i], d[i])
                edge = min(np.linalg.norm(geom[3:6] - geom[0:3]), np.linalg.norm(geom[6:9] - geom[0:3]), np.linalg.norm(geom[6:9] - geom[3:6]))
                print(f"   {tag}: prim {p} triangle, oracle intersect t = {tt}, plane t = {tb!r}, min bary = {mb:.3e}, shortest edge {edge:.3e}, "
                      f"cos = {abs(geom[9:12] @ d[i]) / np.linalg.norm(geom[9:12]):.3e}")
            else:
                print(f"   {tag}: prim {p} sphere c {geom[0:3]} r {geom[3]}, oracle intersect t = {tt}")
        print(f"   t oracle {t0[i]!r}  t gpu {t1[i]!r}")
