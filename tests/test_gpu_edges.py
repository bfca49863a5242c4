"""Edge cases of the per-ray seam and of the render call (GPU, through the C ABI): empty and ragged batches, degenerate rays
(axis-aligned directions give infinite reciprocals and 0 * inf = NaN in the slab test, src/bvh.rs:8-21; zero and NaN directions),
origins on surfaces and inside spheres, one-primitive trees (src/bvh.rs:201-343 with a single leaf), tiny and odd image sizes,
zero-sample render calls.  The f64 instantiation must stay bit-equal to the oracle on every one of them; the production
instantiation must agree away from exact-distance ties."""
import numpy as np
import pytest

from test_host_cpu import _soup_desc

pytestmark = pytest.mark.gpu

CAM = {"image_width": 33, "image_height": 17, "location": [0.0, 1.0, -9.0], "pitch": 0.1, "yaw": 0.0, "roll": 0.0,
       "sensor_width": 0.036, "sensor_height": 0.024, "focal_length": 0.05, "focus_distance": 9.0, "aperture": 8.0}


@pytest.fixture(scope="module")
def pb():
    import paths_b200

    if paths_b200.device_count() < 1:
        pytest.fail("no CUDA device: the product has no CPU path, GPU tests cannot run here")
    return paths_b200


def both(pb, oracle_mod, desc):
    return oracle_mod.Scene(desc), pb.Scene(desc)


def check_f64(oracle_mod, osc, gsc, o, d):
    p0, t0, b0 = osc.trace_closest(o, d, want_bary=True)
    p1, t1, b1 = gsc.find_intersection(o, d, precision=64, want_bary=True)
    assert np.array_equal(p0, p1)
    hit = p0 != oracle_mod.MISS
    # equal_nan: a degenerate ray can "hit" at t = NaN in the reference (NaN fails the `t < 0` rejections of geom.rs:217-231, 273)
    assert np.array_equal(t0[hit], t1[hit], equal_nan=True) and np.array_equal(b0[hit], b1[hit], equal_nan=True)
    return p0, t0, hit


def check_f32(oracle_mod, osc, gsc, o, d, p0, t0, face_grazes=0):
    """ids equal, or the oracle itself finds the other primitive at the same distance (1e-4 relative): a tie.
    `face_grazes`: how many rays may hit in production where the reference misses because the ray lies exactly IN a face plane of
    the primitive's box with a zero direction component -- (box - o) * inv_d is 0 * inf = NaN there and the strict comparisons of
    src/bvh.rs:16-20 cull the box, although the primitive test itself (geom.rs) reports the tangent hit.  The f64 instantiation
    reproduces that; the production slab test (finite reciprocals) does not.  Measure zero for camera and bounce rays."""
    p2, t2 = gsc.find_intersection(o, d, precision=32)
    hit = p0 != oracle_mod.MISS
    same = p0 == p2
    assert np.allclose(t2[same & hit], t0[same & hit], rtol=1e-4, atol=1e-6)
    for i in np.nonzero(~same)[0]:
        if p2[i] != oracle_mod.MISS and face_grazes > 0 and np.any(d[i] == 0.0):
            # the reference culled the nearer primitive's box and reports a miss or something farther
            t_direct = osc.intersect_prim(int(p2[i]), o[i], d[i])
            if t_direct is not None and abs(t_direct - t2[i]) <= 1e-4 * max(abs(t_direct), 1.0) and (p0[i] == oracle_mod.MISS or t_direct < t0[i]):
                face_grazes -= 1
                continue
        assert p0[i] != oracle_mod.MISS and p2[i] != oracle_mod.MISS, f"ray {i}: hit/miss disagreement {p0[i]} vs {p2[i]}"
        t_other = osc.intersect_prim(int(p2[i]), o[i], d[i])
        assert t_other is not None and abs(t_other - t0[i]) <= 1e-4 * max(abs(t0[i]), 1.0), f"ray {i}: {p0[i]} at {t0[i]} vs {p2[i]} at {t_other}"
    return float(np.mean(same))


def test_zero_rays_and_ragged_batches(pb, oracle_mod):
    desc = _soup_desc(np.random.default_rng(1), 300, 4)
    osc, gsc = both(pb, oracle_mod, desc)
    for prec in (32, 64):
        p, t = gsc.find_intersection(np.zeros((0, 3)), np.zeros((0, 3)), precision=prec)
        assert p.shape == (0,) and t.shape == (0,)
    rng = np.random.default_rng(2)
    for n in (1, 31, 32, 33, 127, 129, 4097):  # around warp, block and refill-batch boundaries
        o = rng.uniform(-6, 6, size=(n, 3))
        d = rng.normal(size=(n, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        p0, t0, _ = check_f64(oracle_mod, osc, gsc, o, d)
        check_f32(oracle_mod, osc, gsc, o, d, p0, t0)


def test_axis_aligned_rays(pb, oracle_mod):
    # inv_direction = +-inf on two axes; (box - o) * inf is NaN when the origin sits on a box face (src/bvh.rs:9-14, f64::min/max
    # drop the NaN).  Origins on an integer lattice, boxes from f32-representable vertices: such coincidences do occur.
    desc = _soup_desc(np.random.default_rng(3), 500, 6)
    osc, gsc = both(pb, oracle_mod, desc)
    g = np.arange(-5.0, 5.5, 0.5)
    a, b = np.meshgrid(g, g, indexing="ij")
    a, b = a.ravel(), b.ravel()
    far = np.full_like(a, 9.0)
    rays_o, rays_d = [], []
    for axis in range(3):
        for sign in (1.0, -1.0):
            o = np.zeros((a.size, 3))
            o[:, axis] = -sign * far
            o[:, (axis + 1) % 3] = a
            o[:, (axis + 2) % 3] = b
            d = np.zeros((a.size, 3))
            d[:, axis] = sign
            rays_o.append(o)
            rays_d.append(d)
    o, d = np.concatenate(rays_o), np.concatenate(rays_d)
    p0, t0, hit = check_f64(oracle_mod, osc, gsc, o, d)
    assert hit.mean() > 0.05
    assert check_f32(oracle_mod, osc, gsc, o, d, p0, t0, face_grazes=8) > 0.99  # the lattice lines (x, z) = (+-1, 0), (0, +-1) along +-y are tangent to the light sphere
    # negative zero components take the other sign of infinity
    d2 = np.where(d == 0.0, -0.0, d)
    check_f64(oracle_mod, osc, gsc, o, d2)


def test_degenerate_directions_do_not_hang_and_follow_the_oracle(pb, oracle_mod):
    desc = _soup_desc(np.random.default_rng(4), 200, 3)
    osc, gsc = both(pb, oracle_mod, desc)
    rng = np.random.default_rng(5)
    n = 256
    o = rng.uniform(-3, 3, size=(n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[0::4] = 0.0                       # zero direction: every reciprocal infinite
    d[1::4, 1] = np.nan                 # one NaN component
    d[2::4] *= 1e-30                    # denormal-scale (not unit) direction: t scales by 1e30
    check_f64(oracle_mod, osc, gsc, o, d)
    p2, t2 = gsc.find_intersection(o, d, precision=32)  # must return; the well-formed quarter must still be right
    p0, t0 = osc.trace_closest(o[3::4], d[3::4])
    assert np.mean(p2[3::4] == p0) > 0.98


def test_origins_on_surfaces_and_inside_spheres(pb, oracle_mod):
    desc = _soup_desc(np.random.default_rng(6), 400, 8)
    osc, gsc = both(pb, oracle_mod, desc)
    rng = np.random.default_rng(7)
    o = rng.uniform(-5, 5, size=(20000, 3))
    d = rng.normal(size=o.shape)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    p, t = osc.trace_closest(o, d)
    hit = p != oracle_mod.MISS
    # restart exactly at the hit point (t = 0 self-hits are decided by the sign tests of geom.rs:217-231, 273)
    o2 = (o + d * t[:, None])[hit]
    d2 = rng.normal(size=o2.shape)
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    check_f64(oracle_mod, osc, gsc, o2, d2)
    # origins at sphere centres and just inside: the far root is returned (geom.rs:228-231)
    centres = np.array([ob["center"] for ob in desc["objects"] if ob["shape"] == "Sphere"])
    o3 = np.repeat(centres, 64, axis=0) + rng.normal(scale=0.05, size=(centres.shape[0] * 64, 3))
    d3 = rng.normal(size=o3.shape)
    d3 /= np.linalg.norm(d3, axis=1, keepdims=True)
    p3, t3, hit3 = check_f64(oracle_mod, osc, gsc, o3, d3)
    assert hit3.all()
    assert check_f32(oracle_mod, osc, gsc, o3, d3, p3, t3) > 0.99


@pytest.mark.parametrize("kind", ["triangle", "sphere"])
def test_single_primitive_scene(pb, oracle_mod, kind):
    sky = {"type": "Gradient", "c0": [0.5, 0.7, 0.9], "c1": [0.9, 0.9, 0.9]}
    if kind == "triangle":
        v = np.array([[-2.0, -1.0, 0.5], [2.0, -1.0, -0.5], [0.0, 2.5, 0.25]])
        desc = {"camera": CAM, "models": [{"vertices": v, "faces": np.array([[0, 1, 2]], dtype=np.uint32), "vertex_colours": None, "material": None}],
                "objects": [{"shape": "Mesh", "model": 0, "smooth_normals": True, "translation": [0.0, 0.0, 0.0], "rotation": (0.0, 0.0, 0.0), "scale": 1.0,
                             "material": {"type": "Lambertian", "albedo": {"type": "Rgb", "rgb": [0.7, 0.5, 0.3]}}}], "lights": [], "sky": sky}
    else:
        desc = {"camera": CAM, "models": [], "objects": [{"shape": "Sphere", "center": [0.0, 1.0, 0.0], "radius": 1.5,
                                                          "material": {"type": "Mirror"}}], "lights": [], "sky": sky}
    osc, gsc = both(pb, oracle_mod, desc)
    assert gsc.info["n_prims"] == 1
    o, d, _ = oracle_mod.camera_rays(CAM, 1, 0)
    p0, t0, hit = check_f64(oracle_mod, osc, gsc, o, d)
    assert hit.any() and not hit.all()
    assert check_f32(oracle_mod, osc, gsc, o, d, p0, t0) == 1.0
    # and the image: f64 follows the oracle sample for sample, f32 within rounding of it (no discrete decision flips on this scene's few paths)
    sums, _, _ = osc.render(CAM, 5, 0, 6)
    for prec, tol in ((64, 1e-9), (32, 2e-3)):
        r = pb.Renderer(CAM, gsc, precision=prec, seed=5)
        r.render_samples(6)
        s = r.sums()
        r.shutdown()
        assert np.all(s[..., 3] == 6)
        close = np.isclose(s[..., :3], sums, rtol=tol, atol=tol)
        assert close.mean() > 0.99, (prec, close.mean())


@pytest.mark.parametrize("w,h", [(1, 1), (3, 5), (64, 1), (1, 37), (131, 7)])
def test_tiny_and_odd_images(pb, oracle_mod, w, h):
    cam = dict(CAM, image_width=w, image_height=h)
    desc = _soup_desc(np.random.default_rng(8), 64, 3, with_light=True)
    desc["camera"] = cam
    osc, gsc = both(pb, oracle_mod, desc)
    sums, _, _ = osc.render(cam, 9, 0, 3)
    r = pb.Renderer(cam, gsc, precision=64, seed=9)
    r.render_samples(3)
    s = r.sums()
    img = r.render_rgb24()
    r.shutdown()
    assert s.shape[:2] == (h, w) and np.all(s[..., 3] == 3)
    assert np.isclose(s[..., :3], sums, rtol=1e-9, atol=1e-12).mean() >= 0.99 or w * h < 100
    if w * h < 100:  # too few pixels for a fraction: allow one flipped discrete decision (libm ulp) at most
        assert (~np.isclose(s[..., :3], sums, rtol=1e-9, atol=1e-12).all(axis=-1)).sum() <= 1
    assert np.asarray(img).shape[:2] == (h, w)


def test_zero_sample_render_is_a_no_op(pb, oracle_mod):
    desc = _soup_desc(np.random.default_rng(10), 64, 2)
    desc["camera"] = CAM
    gsc = pb.Scene(desc)
    r = pb.Renderer(CAM, gsc, precision=32, seed=1)
    r.render_samples(0)
    assert r.num_rays_cast() == 0 and np.all(r.sums() == 0)
    r.render_samples(2)
    a = r.sums().copy()
    r.render_samples(0)
    assert np.array_equal(a, r.sums())
    r.reset()
    assert np.all(r.sums() == 0)
    r.shutdown()
