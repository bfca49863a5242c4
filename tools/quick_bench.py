"""Development probe: renders one scene and prints throughput and the per-kernel time split (not the contract bench)."""
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import paths_b200 as pb  # noqa: E402

scene = sys.argv[1] if len(sys.argv) > 1 else "scenes/teapot.yml"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
W = int(sys.argv[3]) if len(sys.argv) > 3 else 1920
H = int(sys.argv[4]) if len(sys.argv) > 4 else 1080
prec = int(sys.argv[5]) if len(sys.argv) > 5 else 32
spass = int(sys.argv[6]) if len(sys.argv) > 6 else 0
import make_assets  # noqa: E402

make_assets.ensure("dragon", "hdri", "doom", "table")
sc = pb.Scene.load(scene, width=W, height=H)
import os  # noqa: E402

if os.environ.get("PATHS_BVH_DEVICE"):
    sc.build_bvh(device=True)
for timed in (False, True):
    r = pb.Renderer(sc.camera, sc, precision=prec, seed=1, time_kernels=timed, samples_per_pass=spass)
    r.render_samples(min(spp, 4))  # warm-up
    r.reset()
    t0 = time.time()
    r.render_samples(spp)
    dt = time.time() - t0
    st = r.stats()
    out = {"scene": scene, "W": W, "H": H, "spp": spp, "precision": prec, "timed": timed, "sec": round(dt, 4),
           "Msamples_s": round(st["samples"] / dt / 1e6, 2), "Mrays_s": round((st["extend_rays"] + st["shadow_rays"]) / dt / 1e6, 2),
           "rays_per_sample": round((st["extend_rays"] + st["shadow_rays"]) / max(st["samples"], 1), 3), "info": sc.info}
    if timed:
        out.update({k: round(v, 3) for k, v in st.items() if k.startswith("ms_")})
        out["launches"] = st["kernel_launches"]
    print(json.dumps(out))
    r.shutdown()
