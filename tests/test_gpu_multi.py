"""The frame combine inside the library (the reference merges its workers' results inside Renderer, src/renderer.rs:93-105).

Tests that need two GPUs are skipped on a one-GPU box (run them with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`;
the log of that run is kept under profiles/).  What they assert: a frame rendered by N ranks and combined over NCCL equals the same
frame rendered by one GPU -- bit for bit for tile shards (disjoint pixels are moved, not added), to f32 rounding for sample shards."""
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
CLI = ROOT / "paths_b200" / "bin" / "paths_cli"


@pytest.fixture(scope="module")
def pb():
    import paths_b200

    if paths_b200.device_count() < 1:
        pytest.fail("no CUDA device: the product has no CPU path")
    return paths_b200


def need(pb, n):
    if pb.device_count() < n:
        pytest.skip(f"needs {n} GPUs, this box has {pb.device_count()}")


def one_gpu(pb, sc, spp, **kw):
    r = pb.Renderer(sc.camera, sc, precision=32, seed=5, **kw)
    r.render_samples(spp)
    s, rgb = r.sums(), r.render_rgb24()
    r.shutdown()
    return s, rgb


def test_group_of_one_device_is_the_plain_renderer(pb):
    sc = pb.Scene.load("scenes/bunny.yml", width=160, height=120)
    ref, rgb = one_gpu(pb, sc, 6)
    g = pb.RendererGroup(sc.camera, sc, 1, precision=32, seed=5)
    g.render_samples(4)
    g.render_samples(2)
    assert np.array_equal(g.sums(), ref) and np.array_equal(g.render_rgb24(), rgb)
    assert g.stats()["samples"] == 160 * 120 * 6
    g.reset()
    assert not g.sums().any()
    g.shutdown()


def test_combine_without_a_communicator_is_refused(pb):
    sc = pb.Scene.load("scenes/spheres_on_plane.yml", width=64, height=48)
    r = pb.Renderer(sc.camera, sc, precision=32, rank=0, world=2)
    with pytest.raises(pb.PathsError) as e:
        r.combine(0)
    assert e.value.status == pb._capi.ERR_INVALID_ARG and "communicator" in str(e.value)
    r.shutdown()
    r = pb.Renderer(sc.camera, sc, precision=32)  # unsharded: the estimator is the frame, combine is a no-op
    r.render_samples(1)
    r.combine(0)
    assert r.sums()[..., 3].min() == 1
    r.shutdown()


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("shard", ["tiles", "samples"])
def test_group_frame_equals_one_gpu(pb, n, shard):
    need(pb, n)
    sc = pb.Scene.load("scenes/bunny.yml", width=640, height=360)  # 240 tiles of 32x32, partial ones at the bottom edge
    spp = 16
    ref, rgb = one_gpu(pb, sc, spp)
    g = pb.RendererGroup(sc.camera, sc, n, precision=32, seed=5, shard_mode=pb.SHARD_TILES if shard == "tiles" else pb.SHARD_SAMPLES)
    g.render_samples(spp // 2)
    g.render_samples(spp // 2)
    got = g.sums()
    assert np.array_equal(got[..., 3], ref[..., 3])
    if shard == "tiles":
        assert np.array_equal(got, ref) and np.array_equal(g.render_rgb24(), rgb)
    else:
        assert np.allclose(got[..., :3], ref[..., :3], rtol=2e-6, atol=1e-6)
    st = g.stats()
    assert st["samples"] == 640 * 360 * spp and st["bad_samples"] == 0
    per = [g.stats(rank=k)["samples"] for k in range(n)]
    assert sum(per) == st["samples"] and min(per) > 0
    g.shutdown()


@pytest.mark.parametrize("shard", ["tiles", "samples"])
def test_process_per_gpu_frame_equals_one_gpu(pb, tmp_path, shard):
    # the torchrun / MPI layout: one process per GPU, unique id handed over out of band, paths_renderer_attach_comm + _combine
    need(pb, 2)
    world, spp, w, h = 2, 5, 400, 300
    procs = [subprocess.Popen([sys.executable, str(ROOT / "tests" / "multi_gpu_worker.py"), str(k), str(world), shard, str(spp), "scenes/teapot.yml",
                               str(w), str(h), str(tmp_path)], cwd=ROOT) for k in range(world)]
    for p in procs:
        assert p.wait(timeout=600) == 0
    got = np.load(tmp_path / "combined.npy")
    sc = pb.Scene.load("scenes/teapot.yml", width=w, height=h)
    # the workers render spp twice: under sample sharding each call's range is split, so the union is samples [0, 2 spp)
    ref, rgb = one_gpu(pb, sc, 2 * spp)
    assert np.array_equal(got[..., 3], ref[..., 3])
    if shard == "tiles":
        assert np.array_equal(got, ref) and np.array_equal(np.load(tmp_path / "rgb.npy"), rgb)
    else:
        assert np.allclose(got[..., :3], ref[..., :3], rtol=2e-6, atol=1e-6)


def test_cli_gpus_writes_the_one_gpu_image(pb, tmp_path):
    need(pb, 2)
    from test_cli import read_ppm

    a, b = tmp_path / "one.ppm", tmp_path / "two.ppm"
    args = ["scenes/bunny.yml", "--spp", "8", "--width", "320", "--height", "200", "--seed", "3", "--precision", "32", "--shard", "tiles"]
    r1 = subprocess.run([str(CLI), *args, "--out", str(a)], cwd=ROOT, capture_output=True, text=True, timeout=300)
    r2 = subprocess.run([str(CLI), *args, "--gpus", "2", "--out", str(b)], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r1.returncode == 0 and r2.returncode == 0, (r1.stderr, r2.stderr)
    assert "2 GPUs" in r2.stdout
    assert np.array_equal(read_ppm(a), read_ppm(b))
