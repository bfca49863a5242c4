"""GPU parity on the large configs of BASELINE.json (dragon.yml, environment.yml, doom.yml).  The reference's /assets are
git-ignored and absent, so seeded synthetic stand-ins of the same size are generated where the YAMLs expect them
(tools/make_assets.py); the scene files themselves are the reference's, unmodified."""
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

    make_assets.ensure("dragon", "hdri", "doom", "table")
    return paths_b200


def relrmse(a, b):
    return float(np.sqrt(np.mean((a - b) ** 2 / (b**2 + 0.01))))


def test_dragon_fixed_rays(pb, oracle_mod):
    # 871 207 primitives: ids and t bit-equal in f64 on camera rays and on incoherent surface rays; f32 ids equal except ties
    osc, cam = oracle_mod.load_scene("scenes/dragon.yml", width=480, height=320)
    gsc = pb.Scene.load("scenes/dragon.yml", width=480, height=320)
    assert gsc.info["n_prims"] == osc.num_prims == 871207
    o, d, _ = oracle_mod.camera_rays(cam, 4, 0)
    p0, t0 = osc.trace_closest(o, d)
    p1, t1 = gsc.find_intersection(o, d, precision=64)
    assert np.array_equal(p0, p1)
    hit = p0 != oracle_mod.MISS
    assert hit.mean() > 0.9 and np.array_equal(t0[hit], t1[hit])
    assert (p0 < 871200).mean() > 0.05  # the mesh is in view
    rng = np.random.default_rng(1)
    sel = np.nonzero(p0 < 871200)[0][:40000]
    o2 = o[sel] + d[sel] * t0[sel, None]
    d2 = rng.normal(size=o2.shape)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    o2 += d2 * 1e-3
    q0, s0 = osc.trace_closest(o2, d2)
    q1, s1 = gsc.find_intersection(o2, d2, precision=64)
    assert np.array_equal(q0, q1) and np.array_equal(s0[q0 != oracle_mod.MISS], s1[q0 != oracle_mod.MISS])
    # production path
    o32, d32 = o.astype(np.float32).astype(np.float64), d.astype(np.float32).astype(np.float64)
    r0, u0, b0 = osc.trace_closest(o32, d32, want_bary=True)
    r1, u1 = gsc.find_intersection(o, d, precision=32)
    same = r0 == r1
    assert same.mean() > 0.995  # 871k sub-pixel triangles: ~1e-3 of the rays pass within f32 position error of an edge
    h = same & (r0 != oracle_mod.MISS)
    assert np.all(np.abs(u1[h] - u0[h]) <= 1e-4 * np.abs(u0[h]))
    # every id difference must be explained by the oracle: the two candidates are a tie within 1e-4 relative (shared edge, the
    # watertight test may pick either neighbour), or the oracle's own hit lies within f32 position error of a triangle edge
    # (silhouette: |o| * 2^-23 against ~0.01-long edges is about 1e-3 in barycentric units)
    unexplained = 0
    for i in np.nonzero(~same)[0]:
        ta = osc.intersect_prim(r0[i], o32[i], d32[i]) if r0[i] != oracle_mod.MISS else None
        tb = osc.intersect_prim(r1[i], o32[i], d32[i]) if r1[i] != oracle_mod.MISS else None
        tie = ta is not None and tb is not None and abs(ta - tb) <= 1e-4 * abs(ta)
        near_edge = r0[i] != oracle_mod.MISS and r0[i] < 871200 and b0[i].min() < 2e-3
        if not (tie or near_edge):
            unexplained += 1
    assert unexplained == 0, f"{unexplained} of {(~same).sum()} id differences are neither ties nor edge grazes"


def test_dragon_image_f32(pb, oracle_mod):
    # NEE to an area light, metallic gloss mesh, five 1e6-radius wall spheres
    osc, cam = oracle_mod.load_scene("scenes/dragon.yml", width=120, height=80)
    gsc = pb.Scene.load("scenes/dragon.yml", width=120, height=80)
    spp = 16
    r = pb.Renderer(cam, gsc, precision=32, seed=4)
    r.render_samples(spp)
    g = r.sums()[..., :3] / spp
    st = r.stats()
    r.shutdown()
    a = osc.render(cam, 4, 0, spp)[0] / spp
    b = osc.render(cam, 5, 0, spp)[0] / spp
    sigma0 = relrmse(a, b)
    assert relrmse(g, a) <= 0.5 * sigma0 and relrmse(g, b) <= 1.15 * sigma0
    assert st["shadow_rays"] > 0 and st["bad_samples"] == 0


@pytest.mark.parametrize("precision", [0, 64])
def test_doom_large_coordinates_follow_oracle_in_f64(pb, oracle_mod, precision):
    # coordinates +-700: AUTO resolves to the mixed instantiation (f64 state, shading and primitive tests over the production tree);
    # both it and the parity instantiation follow the oracle sample for sample.  Per-vertex colours; big area light.
    osc, cam = oracle_mod.load_scene("scenes/doom.yml", width=72, height=48)
    gsc = pb.Scene.load("scenes/doom.yml", width=72, height=48)
    assert gsc.info["auto_precision"] == pb.PRECISION_MIXED == 48
    r = pb.Renderer(cam, gsc, seed=6, precision=precision)
    _, _, eb = r.device_sums()
    assert eb == 8
    r.render_samples(3)
    s = r.sums()
    st = r.stats()
    r.shutdown()
    sums, counts, ctr = osc.render(cam, 6, 0, 3)
    close = np.all(np.abs(s[..., :3] - sums) <= 1e-9 * (np.abs(sums) + 1e-3), axis=-1)
    assert close.mean() >= 0.995 and sums.max() > 0
    assert st["shadow_rays"] > 0 and st["bad_samples"] == 0


def test_mixed_fixed_rays_are_bit_exact_away_from_ties(pb, oracle_mod):
    # production tree + the reference's f64 primitive tests: t and barycentrics bit-equal, ids equal unless two primitives tie exactly
    for name, kw in (("doom", dict(width=240, height=160)), ("dragon", dict(width=240, height=160)), ("teapot", dict(width=240, height=160))):
        osc, cam = oracle_mod.load_scene(f"scenes/{name}.yml", **kw)
        gsc = pb.Scene.load(f"scenes/{name}.yml", **kw)
        o, d, _ = oracle_mod.camera_rays(cam, 4, 0)
        p0, t0, b0 = osc.trace_closest(o, d, want_bary=True)
        p1, t1, b1 = gsc.find_intersection(o, d, precision=pb.PRECISION_MIXED, want_bary=True)
        hit = p0 != oracle_mod.MISS
        assert np.array_equal(hit, p1 != oracle_mod.MISS), name
        assert np.array_equal(t0[hit], t1[hit]), name  # the closest distance itself never depends on the tree
        same = p0 == p1
        assert same.mean() > 0.9999, name
        assert np.array_equal(b0[same & hit], b1[same & hit]), name
        # incoherent rays off the surfaces
        rng = np.random.default_rng(2)
        sel = np.nonzero(hit)[0][:30000]
        o2 = o[sel] + d[sel] * t0[sel, None]
        d2 = rng.normal(size=o2.shape)
        d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
        o2 += d2 * 1e-3
        q0, s0 = osc.trace_closest(o2, d2)
        q1, s1 = gsc.find_intersection(o2, d2, precision=pb.PRECISION_MIXED)
        h2 = q0 != oracle_mod.MISS
        assert np.array_equal(h2, q1 != oracle_mod.MISS) and np.array_equal(s0[h2], s1[h2]), name
        assert (q0 == q1).mean() > 0.9999, name


def test_environment_hdri_and_auto_material(pb, oracle_mod):
    # HDRI lookup on a miss (scene.rs:95-111), `material: Auto` from the OBJ's MTL, smooth-normal dragon
    osc, cam = oracle_mod.load_scene("scenes/environment.yml", width=96, height=64)
    gsc = pb.Scene.load("scenes/environment.yml", width=96, height=64)
    r = pb.Renderer(cam, gsc, precision=64, seed=8)
    r.render_samples(2)
    s = r.sums()
    r.shutdown()
    sums, _, _ = osc.render(cam, 8, 0, 2)
    close = np.all(np.abs(s[..., :3] - sums) <= 1e-9 * (np.abs(sums) + 1e-3), axis=-1)
    assert close.mean() >= 0.995
    r = pb.Renderer(cam, gsc, precision=32, seed=8)
    r.render_samples(16)
    g = r.sums()[..., :3] / 16
    r.shutdown()
    a = osc.render(cam, 8, 0, 16)[0] / 16
    b = osc.render(cam, 9, 0, 16)[0] / 16
    assert relrmse(g, a) <= 0.5 * relrmse(a, b)


def test_full_size_properties_teapot_1080p(pb):
    # BASELINE's headline size, through properties that need no oracle run: the estimator is a pure function of (seed, pixel, sample),
    # whatever the launch order, pass size, number of pass contexts, shard layout or call pattern
    sc = pb.Scene.load("scenes/teapot.yml", width=1920, height=1080)

    def sums(spp_calls, **kw):
        r = pb.Renderer(sc.camera, sc, precision=32, seed=21, **kw)
        for n in spp_calls:
            r.render_samples(n)
        s = r.sums()
        st = r.stats()
        r.shutdown()
        return s, st

    a, st = sums([20])
    assert st["samples"] == 20 * 1920 * 1080 and st["bad_samples"] == 0
    assert np.all(a[..., 3] == 20) and np.isfinite(a).all() and a[..., :3].min() >= 0
    b, _ = sums([20])
    assert np.array_equal(a, b)                      # run to run
    c, _ = sums([3, 1, 16])
    assert np.array_equal(a, c)                      # additivity over calls: render(3) + render(1) + render(16) == render(20)
    d, _ = sums([20], samples_per_pass=3)
    assert np.array_equal(a, d)                      # pass size (7 passes instead of 2)
    # shards: 3 ranks by tiles are disjoint and bit-identical; 3 ranks by sample ranges sum to the same image up to f32 rounding
    tiles = [sums([20], rank=k, world=3, shard_mode=pb.SHARD_TILES)[0] for k in range(3)]
    assert np.array_equal(sum(tiles), a) and np.all(sum((t[..., 3] > 0).astype(int) for t in tiles) == 1)
    parts = [sums([20], rank=k, world=3, shard_mode=pb.SHARD_SAMPLES)[0] for k in range(3)]
    tot = sum(parts)
    assert np.array_equal(tot[..., 3], a[..., 3]) and np.allclose(tot[..., :3], a[..., :3], rtol=2e-6, atol=1e-6)
    # tonemap: monotone in the mean, saturating (colour.rs:62-74)
    r = pb.Renderer(sc.camera, sc, precision=32, seed=21)
    r.render_samples(4)
    img, rgb = r.render().pixels, r.render_rgb24()
    r.shutdown()
    m = img[..., 0].ravel()
    order = np.argsort(m, kind="stable")
    assert np.all(np.diff(rgb[..., 0].ravel()[order].astype(int)) >= 0)
    assert np.all(rgb[img >= 1.0] == 255) and np.all(rgb[img <= 0.0] == 0)


@pytest.mark.parametrize("name,ref_spp", [("teapot", 2048), ("dragon", 1024)])
def test_headline_convergence(pb, oracle_mod, name, ref_spp):
    # SURVEY 8(d), Monte Carlo tolerance of the headline scenes, at reduced resolution so that the CPU reference finishes in under a
    # minute.  Gate (ii): at the config's 256 spp the GPU image is as close to a many-spp oracle image as a second oracle run is
    # (sigma0); the error falls as samples are added; the image mean is unbiased.  teapot also meets the absolute figure: the error
    # follows 1/sqrt(N) (0.224, 0.113, 0.058 at 16, 64, 256 spp), so relRMSE 0.05 is this estimator's noise at ~350 spp and is asserted
    # at 512.  dragon (small bright area light seen through glossy bounces) is firefly-dominated in linear radiance for the reference
    # estimator itself, so it is compared display-referred: radiance clamped to [0, 1] like Colour::to_bytes does (colour.rs:62-74).
    w, h = 160, 90
    osc, cam = oracle_mod.load_scene(f"scenes/{name}.yml", width=w, height=h)
    gsc = pb.Scene.load(f"scenes/{name}.yml", width=w, height=h)
    clamp = (lambda im: np.clip(im, 0.0, 1.0)) if name == "dragon" else (lambda im: im)
    sums, _, _ = osc.render(cam, 77, 0, ref_spp)
    ref = clamp(sums / ref_spp)
    sums_b, _, _ = osc.render(cam, 78, 0, 256)
    sigma0 = relrmse(clamp(sums_b / 256), ref)
    r = pb.Renderer(cam, gsc, precision=32, seed=5)
    errs, done = [], 0
    for target in (16, 64, 256, 512):
        r.render_samples(target - done)
        done = target
        s = r.sums()
        errs.append(relrmse(clamp(s[..., :3] / s[..., 3:]), ref))
    st = r.stats()
    r.shutdown()
    assert st["bad_samples"] == 0
    assert errs[2] <= 1.15 * sigma0, (errs, sigma0)
    assert errs[0] > errs[1] > errs[2] > errs[3], errs
    g = s[..., :3] / s[..., 3:]
    assert abs(clamp(g).mean() - ref.mean()) <= 0.01 * ref.mean()
    if name == "teapot":
        assert errs[3] <= 0.05, errs
        assert 0.42 <= errs[1] / errs[0] <= 0.60 and 0.42 <= errs[2] / errs[1] <= 0.65, errs
