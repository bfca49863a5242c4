"""Development probe: sha256 of the estimator sums of a short render -- equal hashes across library builds (PATHS_B200_LIB) mean that every
ray got the same answer.   python tools/image_hash.py scenes/doom.yml 4 480 270 48"""
import hashlib
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import paths_b200 as pb  # noqa: E402
import make_assets  # noqa: E402

scene = sys.argv[1]
spp, W, H, prec = (int(x) for x in sys.argv[2:6])
make_assets.ensure("dragon", "hdri", "doom", "table")
sc = pb.Scene.load(scene, width=W, height=H)
r = pb.Renderer(sc.camera, sc, precision=prec, seed=5)
r.render_samples(spp)
st = r.stats()
h = hashlib.sha256(r.sums().tobytes()).hexdigest()[:16]
print(f"{scene} {W}x{H} {spp} spp precision {prec}: sums sha256 {h} extend_rays {st['extend_rays']} shadow_rays {st['shadow_rays']} bad {st['bad_samples']}")
r.shutdown()
