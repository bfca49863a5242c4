"""One rank of a process-per-GPU frame (spawned by tests/test_gpu_multi.py): renders its shard through the C ABI, joins the frame's NCCL
communicator with the unique id rank 0 published in a file, combines on rank 0 and, on rank 0, saves the combined estimator."""
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    rank, world, shard, spp, scene, w, h, out_dir = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4]), sys.argv[5], int(sys.argv[6]), int(sys.argv[7]), Path(sys.argv[8])
    import paths_b200 as pb

    id_file = out_dir / "nccl_id.bin"
    if rank == 0:
        uid = pb.comm_unique_id()
        tmp = out_dir / "nccl_id.tmp"
        tmp.write_bytes(uid)
        tmp.rename(id_file)
    else:
        for _ in range(600):
            if id_file.exists():
                break
            time.sleep(0.1)
        uid = id_file.read_bytes()
    sc = pb.Scene.load(scene, width=w, height=h)
    r = pb.Renderer(sc.camera, sc, device=rank, precision=32, seed=5, rank=rank, world=world,
                    shard_mode=pb.SHARD_TILES if shard == "tiles" else pb.SHARD_SAMPLES)
    r.attach_comm(uid)
    # two increments with a combine after each: the combined frame must track the cumulative estimator of every rank
    r.render_samples(spp)
    r.combine(0)
    r.render_samples(spp)
    r.combine(0)
    if rank == 0:
        np.save(out_dir / "combined.npy", r.sums())
        np.save(out_dir / "rgb.npy", r.render_rgb24())
    r.shutdown()


if __name__ == "__main__":
    main()
