"""Every kernel family once at tiny sizes (f32 / f64 / mixed, host and device tree builders, preview, column sweeps, whole frames).
Written as the workload for compute-sanitizer, which is closed on this pool; it now runs under the debug-assert build instead
(tools/run_asserts.sh sets PATHS_B200_LIB to the build made with -DPATHS_DEBUG_ASSERTS and checks its counters afterwards)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np  # noqa: E402

import paths_b200 as pb  # noqa: E402

for name, prec, dev in (("bunny", 32, False), ("bunny", 64, False), ("bunny", 48, True), ("teapot", 32, True)):
    sc = pb.Scene.load(f"scenes/{name}.yml", width=48, height=32)
    if dev:
        sc.build_bvh(device=True)
        assert sc.check_bvh() == 0
    r = pb.Renderer(sc.camera, sc, precision=prec, seed=1)
    r.request_preview()
    r.render_columns(48)
    r.render_samples(2)
    img = r.render_rgb24()
    o, d, _ = pb.camera_rays(sc.camera, 1, 0)
    sc.find_intersection(o, d, precision=prec)
    st = r.stats()
    r.shutdown()
    print(name, prec, dev, st["samples"], int(img.sum()))
sc = pb.Scene.stress(200, seed=1)
r = pb.Renderer(sc.camera, sc, precision=32, seed=1, samples_per_pass=2)
r.render_samples(2)
print("stress", r.stats()["samples"])
r.shutdown()
