"""World-size-2 test of the multi-rank host path on CPU (gloo): each rank renders its shard with the oracle, partial
estimator buffers are combined with paths_b200.shard.combine, and rank 0 checks the frame against the unsharded render.
Tile shards must be bit-identical (disjoint pixels), sample shards equal up to summation order."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

W, H, SPP, SEED = 96, 80, 6, 17


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    from paths_b200 import shard

    osc, cam = oracle.load_scene("scenes/spheres_on_plane.yml", width=W, height=H)
    results = {}
    # ---- sample-range sharding
    lo, hi = shard.sample_slice(0, SPP, rank, world)
    sums, counts, _ = osc.render(cam, SEED, lo, hi, threads=2)
    buf = torch.from_numpy(np.concatenate([sums, counts[..., None].astype(np.float64)], axis=-1).copy())
    got = shard.combine(buf)
    if rank == 0:
        results["samples"] = got.numpy().copy()
    # ---- tile sharding (the oracle's owns_pixel is the same rule as shard.tile_owner / the CUDA renderer)
    sums, counts, _ = osc.render(cam, SEED, 0, SPP, shard_mode=1, rank=rank, world=world, threads=2)
    own = shard.owned_pixels(W, H, rank, world)
    mask = np.zeros(W * H, dtype=bool)
    mask[own] = True
    assert np.array_equal(counts.reshape(-1) > 0, mask), "oracle tile ownership differs from paths_b200.shard"
    buf = torch.from_numpy(np.concatenate([sums, counts[..., None].astype(np.float64)], axis=-1).copy())
    got = shard.combine(buf)
    if rank == 0:
        results["tiles"] = got.numpy().copy()
        np.savez(out_path, **results)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_combine_to_the_full_frame(oracle_mod, tmp_path):
    out = str(tmp_path / "combined.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    osc, cam = oracle_mod.load_scene("scenes/spheres_on_plane.yml", width=W, height=H)
    sums, counts, _ = osc.render(cam, SEED, 0, SPP)
    ref = np.concatenate([sums, counts[..., None].astype(np.float64)], axis=-1)
    assert np.array_equal(got["tiles"], ref)
    assert np.array_equal(got["samples"][..., 3], ref[..., 3])
    assert np.allclose(got["samples"][..., :3], ref[..., :3], rtol=1e-12, atol=1e-12)


def test_sample_slices_partition_the_range():
    from paths_b200 import shard

    for world in (1, 2, 3, 4, 8):
        for spp in (1, 5, 8, 64, 100):
            seen = []
            for r in range(world):
                lo, hi = shard.sample_slice(10, spp, r, world)
                seen += list(range(lo, hi))
            assert seen == list(range(10, 10 + spp))
    tiles = {shard.tile_owner(x, y, 200, 3) for x in range(0, 200, 32) for y in range(0, 136, 32)}
    assert tiles == {0, 1, 2}
    assert sum(len(shard.owned_pixels(200, 136, r, 3)) for r in range(3)) == 200 * 136
