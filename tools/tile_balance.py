"""Development probe (one GPU): per-rank device time of a tile-sharded frame, every rank rendered in turn on the same GPU --
the load imbalance a tile size / dealing rule would show on `world` GPUs.   python tools/tile_balance.py environment 8 [spp]"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))
import make_assets  # noqa: E402
import paths_b200 as pb  # noqa: E402

scene = sys.argv[1] if len(sys.argv) > 1 else "environment"
world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 64
make_assets.ensure("dragon", "hdri", "table", "doom")
sc = pb.Scene.load(f"scenes/{scene}.yml", width=1920, height=1080)
ms, rays = [], []
for rank in range(world):
    r = pb.Renderer(sc.camera, sc, seed=1, rank=rank, world=world, shard_mode=pb.SHARD_TILES)
    r.render_samples(spp)  # warm-up
    r.reset()
    r.render_samples(spp)
    st = r.stats()
    ms.append(st["ms_device"])
    rays.append((st["extend_rays"] + st["shadow_rays"]) / 1e6)
    r.shutdown()
print(scene, "world", world, "lib", pb._capi.LIB_PATH.parent.parent.parent.name, "rank ms", [round(x, 2) for x in ms], "max/mean %.3f" % (max(ms) / (sum(ms) / len(ms))),
      "Mrays", [round(x, 2) for x in rays], "max/mean %.3f" % (max(rays) / (sum(rays) / len(rays))))
