"""The rows SURVEY 8(f) marks "next", through the C ABI on the GPU: the seeded stress scene (src/stress.rs), the interactive front
half of Renderer (preview grid + centre-out column sweeps, src/renderer.rs:143-192, src/pixels.rs:53-79) and checkpoint/resume."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pb():
    import paths_b200

    if paths_b200.device_count() < 1:
        pytest.fail("no CUDA device: the product has no CPU path")
    return paths_b200


def test_stress_scene_matches_the_oracle_generator(pb, oracle_mod):
    from oracle import scene_io

    desc = scene_io.stress_scene(500, seed=7)
    osc = oracle_mod.Scene(desc)
    gsc = pb.Scene.stress(500, seed=7)
    assert gsc.info["n_prims"] == osc.num_prims == 500 and gsc.info["n_spheres"] == 500
    for p in (0, 17, 499):  # the same spheres, bit for bit
        is_tri, g = gsc.prim_geometry(p)
        o = desc["objects"][p]
        assert not is_tri and list(g[:3]) == o["center"] and g[3] == o["radius"]
    cam = dict(desc["camera"], image_width=240, image_height=160)
    o, d, _ = oracle_mod.camera_rays(cam, 5, 0)
    p0, t0 = osc.trace_closest(o, d)
    p1, t1 = gsc.find_intersection(o, d, precision=64)
    assert np.array_equal(p0, p1) and np.array_equal(t0[p0 != oracle_mod.MISS], t1[p0 != oracle_mod.MISS])
    p2, t2 = gsc.find_intersection(o, d, precision=32)
    assert (p0 == p2).mean() > 0.999
    # image: the f64 instantiation follows the oracle sample for sample (Gloss r0 in [1, 3], mirrors, no lights)
    r = pb.Renderer(cam, gsc, precision=64, seed=3)
    r.render_samples(2)
    s = r.sums()
    r.shutdown()
    sums, counts, _ = osc.render(cam, 3, 0, 2)
    close = np.all(np.abs(s[..., :3] - sums) <= 1e-9 * np.abs(sums) + 1e-12, axis=-1)
    assert close.mean() > 0.995


def _centre_out(w):
    cols = []
    for n in range(w):
        x = (w + n) // 2 if n % 2 == 0 else (w - n) // 2
        if w % 2 == 1 and n == 1:
            x = 0
        cols.append(x)
    return cols


@pytest.mark.parametrize("precision", [32, 64])
def test_column_sweeps_equal_whole_frame_rendering(pb, precision):
    sc = pb.Scene.load("scenes/bunny.yml", width=96, height=64)
    a = pb.Renderer(sc.camera, sc, precision=precision, seed=4)
    a.render_columns(10)
    cnt = a.sums()[..., 3]
    want = np.zeros((64, 96))
    want[:, _centre_out(96)[:10]] = 1
    assert np.array_equal(cnt, want) and a.progress() == (0, 10)  # renderer.rs:166-170: centre first, alternating sides
    with pytest.raises(pb.PathsError):
        a.render_samples(1)  # not in the middle of a sweep
    a.render_columns(86)            # completes the (1,1) sweep
    assert a.progress() == (1, 0)
    a.render_columns(96 + 40)       # one whole (5,5) sweep and part of the next
    assert a.progress() == (26, 40)
    a.render_columns(56)
    assert a.progress() == (51, 0)
    b = pb.Renderer(sc.camera, sc, precision=precision, seed=4)
    b.render_samples(51)
    assert np.array_equal(a.sums(), b.sums())  # same samples, accumulated in the same order
    a.shutdown()
    b.shutdown()


def test_odd_width_sweep_covers_every_column_once(pb):
    sc = pb.Scene.load("scenes/spheres_on_plane.yml", width=45, height=20)
    r = pb.Renderer(sc.camera, sc, precision=32, seed=1)
    r.render_columns(45)
    assert np.all(r.sums()[..., 3] == 1) and sorted(_centre_out(45)) == list(range(45))
    r.shutdown()


def _estimator_render(sums, grid=6):
    """Estimator::render (src/pixels.rs:53-79) restated with numpy: unsampled pixels show their preview-grid pixel."""
    h, w, _ = sums.shape
    out = np.zeros((h, w, 3))
    for y in range(h):
        for x in range(w):
            s = sums[y, x]
            if s[3] == 0:
                s = sums[y - y % grid, x - x % grid]
            out[y, x] = s[:3] / s[3] if s[3] > 0 else 0.0
    return out


def test_preview_grid(pb, oracle_mod):
    sc = pb.Scene.load("scenes/teapot.yml", width=100, height=62)
    r = pb.Renderer(sc.camera, sc, precision=64, seed=2)
    r.request_preview()
    s = r.sums()
    want = np.zeros((62, 100))
    want[::6, ::6] = 1
    assert np.array_equal(s[..., 3], want)  # renderer.rs:152-164
    img = r.render().pixels
    ref = _estimator_render(s)
    assert np.allclose(img, ref.astype(np.float32), rtol=1e-6, atol=0)
    assert np.array_equal(r.render_rgb24(), oracle_mod.to_bytes(ref.reshape(-1, 3)).reshape(62, 100, 3))
    # regular samples then replace the fill, column by column
    r.render_columns(100)
    s2 = r.sums()
    assert np.array_equal(s2[..., 3], want + 1)
    assert np.allclose(r.render().pixels, (s2[..., :3] / s2[..., 3:]).astype(np.float32), rtol=1e-6)
    # a camera move starts a new epoch: estimator and schedule are back at zero
    r.reset()
    assert r.progress() == (0, 0) and not r.sums().any()
    r.shutdown()


def test_checkpoint_resume_is_bit_identical(pb, tmp_path):
    sc = pb.Scene.load("scenes/bunny.yml", width=120, height=80)
    ck = tmp_path / "bunny.ck"
    a = pb.Renderer(sc.camera, sc, precision=32, seed=11)
    a.render_samples(8)
    a.save_checkpoint(ck)
    a.render_samples(8)
    b = pb.Renderer(sc.camera, sc, precision=32, seed=11)
    b.load_checkpoint(ck)
    assert b.progress() == (8, 0)
    b.render_samples(8)
    assert np.array_equal(a.sums(), b.sums())
    c = pb.Renderer(sc.camera, sc, precision=32, seed=12)
    with pytest.raises(pb.PathsError):
        c.load_checkpoint(ck)  # another seed would repeat samples
    with pytest.raises(pb.PathsError):
        c.load_checkpoint(tmp_path / "missing.ck")
    # the header also records what else decides the estimator: other bounce limits, extensions, precision or shard layouts are refused
    for kw in (dict(max_bounces=3), dict(extensions=pb.EXT_NEE_ALL_LIGHTS), dict(precision=64), dict(rank=0, world=2)):
        opts = dict(precision=32, seed=11)
        opts.update(kw)
        d = pb.Renderer(sc.camera, sc, **opts)
        with pytest.raises(pb.PathsError) as e:
            d.load_checkpoint(ck)
        assert "other render options" in str(e.value)
        d.shutdown()
    for r in (a, b, c):
        r.shutdown()
