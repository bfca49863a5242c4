"""The program `tools/capture_roofline.sh` runs under ncu: ONE wavefront pass of a bench scene (the bench's pass size: 16 spp at 1080p),
single pass context so that launches do not overlap, after one warm-up pass.  Prints the run's own counters as one JSON line."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))
import paths_b200 as pb  # noqa: E402

scene = sys.argv[1] if len(sys.argv) > 1 else "teapot"
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 1
if scene in ("dragon", "environment", "doom"):
    import make_assets

    make_assets.ensure({"dragon": "dragon", "environment": "dragon", "doom": "doom"}[scene], *(("hdri", "table") if scene == "environment" else ()))
sc = pb.Scene.load(f"scenes/{scene}.yml", width=1920, height=1080)
r = pb.Renderer(sc.camera, sc, precision=0, seed=1, time_kernels=True)  # time_kernels => one pass context, launches serialised
r.render_samples(16)  # warm-up pass (the capture skips its launches)
st0 = r.stats()
r.render_samples(16 * passes)
st1 = r.stats()
d = {k: st1[k] - st0[k] for k in st1 if k not in ("extend_kernel", "camera_kernel", "shade_kernel", "pass_contexts", "stack_overflows")}
d.update({k: st1[k] for k in ("extend_kernel", "camera_kernel", "shade_kernel", "pass_contexts", "stack_overflows")})
d["scene"] = scene
d["launches_per_pass"] = (st1["kernel_launches"] - st0["kernel_launches"]) // passes
d["warmup_launches"] = st0["kernel_launches"]
print(json.dumps(d))
r.shutdown()
