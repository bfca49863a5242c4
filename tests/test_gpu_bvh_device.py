"""The production tree built on the GPU (bvh_device.cu) against the oracle and against the host SAH tree.

Closest hit is independent of the tree's shape except for exact-t ties, so the device-built morton tree must return the same
primitives as the reference algorithm on fixed rays and the same images as the host-built tree."""
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def pb():
    import paths_b200

    if paths_b200.device_count() < 1:
        pytest.fail("no CUDA device: the product has no CPU path")
    sys.path.insert(0, str(ROOT / "tools"))
    import make_assets

    make_assets.ensure("dragon")
    return paths_b200


def _explain(osc, oracle_mod, o32, d32, r0, r1, b0=None, n_mesh=None):
    unexplained = 0
    for i in np.nonzero(r0 != r1)[0]:
        ta = osc.intersect_prim(r0[i], o32[i], d32[i]) if r0[i] != oracle_mod.MISS else None
        tb = osc.intersect_prim(r1[i], o32[i], d32[i]) if r1[i] != oracle_mod.MISS else None
        tie = ta is not None and tb is not None and abs(ta - tb) <= 1e-4 * abs(ta)
        near_edge = b0 is not None and r0[i] != oracle_mod.MISS and r0[i] < n_mesh and b0[i].min() < 2e-3
        if not (tie or near_edge):
            unexplained += 1
    return unexplained


@pytest.mark.parametrize("name", ["teapot", "bunny"])
def test_device_tree_small_meshes(pb, oracle_mod, name):
    osc, cam = oracle_mod.load_scene(f"scenes/{name}.yml", width=360, height=240)
    gsc = pb.Scene.load(f"scenes/{name}.yml", width=360, height=240)
    gsc.build_bvh(device=True)
    assert gsc.check_bvh() == 0
    info = gsc.info
    assert info["bvh_builder"] == 1 and info["device_nodes4"] > 0 and info["build_ms_device"] > 0
    geo = [gsc.prim_geometry(p) for p in range(info["n_prims"])]
    n_big = sum(1 for is_tri, g in geo if not is_tri and g[3] > 1e4)
    assert info["giant_prims"] == n_big  # the 1e6-radius plane spheres sit above the morton root
    o, d, _ = oracle_mod.camera_rays(cam, 11, 0)
    o32, d32 = o.astype(np.float32).astype(np.float64), d.astype(np.float32).astype(np.float64)
    p0, t0 = osc.trace_closest(o32, d32)
    p1, t1 = gsc.find_intersection(o, d, precision=32)
    same = p0 == p1
    hit = same & (p0 != oracle_mod.MISS)
    assert np.all(np.abs(t1[hit] - t0[hit]) <= 1e-4 * np.abs(t0[hit]))
    assert (~same).sum() <= 2e-4 * len(p0)
    assert _explain(osc, oracle_mod, o32, d32, p0, p1) == 0
    # same answers as the host-built tree on incoherent rays, bit for bit away from ties
    host = pb.Scene.load(f"scenes/{name}.yml", width=360, height=240)
    rng = np.random.default_rng(5)
    sel = np.nonzero(p0 != oracle_mod.MISS)[0][:50000]
    o2 = o[sel] + d[sel] * t0[sel, None]
    d2 = rng.normal(size=o2.shape)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    o2 += d2 * 1e-3
    qa, ta = host.find_intersection(o2, d2, precision=32)
    qb, tb = gsc.find_intersection(o2, d2, precision=32)
    eq = qa == qb
    assert eq.mean() > 0.9995
    assert np.array_equal(ta[eq], tb[eq])


def test_device_tree_dragon(pb, oracle_mod):
    osc, cam = oracle_mod.load_scene("scenes/dragon.yml", width=480, height=320)
    gsc = pb.Scene.load("scenes/dragon.yml", width=480, height=320)
    gsc.build_bvh(device=True)
    assert gsc.check_bvh() == 0
    info = gsc.info
    assert info["bvh_builder"] == 1 and info["giant_prims"] == 6
    # build time: the first build of a process also pays for loading the builder's kernels, so a second staging is timed
    # and the best of three (the event-timed region includes the scratch cudaMalloc, whose cost depends on what the process freed before)
    host_ms = gsc.build_bvh()["build_ms_fast"]
    dev_ms = []
    for _ in range(3):
        gsc.build_bvh()
        gsc.build_bvh(device=True)
        assert gsc.check_bvh() == 0  # stages the scene: this is where the device build runs
        dev_ms.append(gsc.info["build_ms_device"])
    assert min(dev_ms) < 100.0 and min(dev_ms) < 0.5 * host_ms, (dev_ms, host_ms)  # the host SAH build takes 0.5 - 0.9 s
    o, d, _ = oracle_mod.camera_rays(cam, 4, 0)
    o32, d32 = o.astype(np.float32).astype(np.float64), d.astype(np.float32).astype(np.float64)
    r0, u0, b0 = osc.trace_closest(o32, d32, want_bary=True)
    r1, u1 = gsc.find_intersection(o, d, precision=32)
    same = r0 == r1
    assert same.mean() > 0.995
    h = same & (r0 != oracle_mod.MISS)
    assert np.all(np.abs(u1[h] - u0[h]) <= 1e-4 * np.abs(u0[h]))
    assert _explain(osc, oracle_mod, o32, d32, r0, r1, b0, 871200) == 0


def test_device_tree_renders_the_same_image(pb):
    # four-wide kernels over the device-built tree: the image equals the host-tree image except where a tie flips a path
    imgs = []
    for dev in (False, True):
        sc = pb.Scene.load("scenes/dragon.yml", width=240, height=160)
        if dev:
            sc.build_bvh(device=True)
        r = pb.Renderer(sc.camera, sc, precision=32, seed=9)
        r.render_samples(8)
        imgs.append(r.sums()[..., :3] / 8)
        st = r.stats()
        assert st["bad_samples"] == 0
        r.shutdown()
    a, b = imgs
    same = np.all(a == b, axis=-1)
    assert same.mean() > 0.99
    assert abs(a.mean() - b.mean()) < 0.01 * a.mean()


def test_builder_switch_and_small_scene_fallback(pb, oracle_mod):
    sc = pb.Scene.load("scenes/spheres_on_plane.yml", width=96, height=64)
    sc.build_bvh(device=True)  # 4 primitives: the device builder declines, the host tree is used
    osc, cam = oracle_mod.load_scene("scenes/spheres_on_plane.yml", width=96, height=64)
    o, d, _ = oracle_mod.camera_rays(cam, 3, 0)
    p, t = sc.find_intersection(o, d, precision=32)
    assert sc.info["bvh_builder"] == 0 and sc.check_bvh() == 0
    p0, _ = osc.trace_closest(o, d)
    assert (p == p0).mean() > 0.999
    sc2 = pb.Scene.load("scenes/teapot.yml", width=96, height=64)
    _, cam2 = oracle_mod.load_scene("scenes/teapot.yml", width=96, height=64)
    o, d, _ = oracle_mod.camera_rays(cam2, 3, 0)
    pa, ta = sc2.find_intersection(o, d, precision=32)
    sc2.build_bvh(device=True)
    pb_, tb = sc2.find_intersection(o, d, precision=32)
    assert sc2.info["bvh_builder"] == 1
    sc2.build_bvh()  # back to the host SAH tree
    pc, tc = sc2.find_intersection(o, d, precision=32)
    assert np.array_equal(pa, pc) and np.array_equal(ta, tc)
    assert (pa == pb_).mean() > 0.9995
