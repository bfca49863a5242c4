"""Estimator extensions that have no reference behaviour (include/paths_b200.h PATHS_EXT_*): environment NEE
through a CDF table, microfacet sampling where Material::sample panics, GGX, all-lights NEE weighting.  There is nothing in the
reference to be in parity with, so each is checked for UNBIASEDNESS against an independent estimator of the same integral,
and the default (extensions = 0) is checked to still follow the oracle, including where the reference would panic."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CAM = {"image_width": 64, "image_height": 48, "location": [0.0, 2.0, -7.0], "pitch": 0.2, "yaw": 0.0, "roll": 0.0,
       "sensor_width": 0.036, "sensor_height": 0.024, "focal_length": 0.05, "focus_distance": 7.0, "aperture": 16.0}


def lam(rgb):
    return {"type": "Lambertian", "albedo": {"type": "Rgb", "rgb": rgb}}


def sphere(c, r, mat):
    return {"shape": "Sphere", "center": c, "radius": r, "material": mat}


@pytest.fixture(scope="module")
def pb():
    import paths_b200

    if paths_b200.device_count() < 1:
        pytest.fail("no CUDA device")
    return paths_b200


def render_mean(pb, desc, spp, seed=1, precision=32, **kw):
    sc = pb.Scene(desc)
    r = pb.Renderer(CAM, sc, precision=precision, seed=seed, **kw)
    r.render_samples(spp)
    s = r.sums()
    st = r.stats()
    r.shutdown()
    return s[..., :3] / s[..., 3:4], st


def sun_sky(w=256, h=128, sun_radius_deg=5.0, sun=400.0):
    ys, xs = np.mgrid[0:h, 0:w]
    tex = np.full((h, w, 3), 0.08, dtype=np.float32)
    tex[..., 2] = 0.12
    # the reference looks the map up with the NEGATED ray direction (trace.rs:21): texels in the lower half light the scene from above
    d2 = ((xs - 0.62 * w) * (2 * np.pi / w)) ** 2 + ((ys - 0.28 * h) * (np.pi / h)) ** 2
    tex[d2 < np.radians(sun_radius_deg) ** 2] = np.array([sun, sun * 0.9, sun * 0.7], dtype=np.float32)
    return {"type": "Hdri", "c0": [0, 0, 0], "c1": [0, 0, 0], "width": w, "height": h, "texels": tex}


def test_environment_nee_matches_the_analytic_direct_term(pb):
    # Lambertian ground under a sun disk, direct light only (max_bounces = 0): the NEE estimator must converge to
    # rho / pi * sum_texels L_t max(0, n . w_t) dOmega_t, computed here from the texture itself.
    sky = sun_sky()
    rho = np.array([0.7, 0.6, 0.5])
    cam = dict(CAM, location=[0.0, 3.0, -5.0], pitch=0.9)
    desc = {"camera": cam, "models": [], "lights": [], "sky": sky, "objects": [sphere([0.0, -1e5, 0.0], 1e5, lam(list(rho)))]}
    sc = pb.Scene(desc)
    r = pb.Renderer(cam, sc, precision=32, seed=2, max_bounces=0, extensions=pb.EXT_ENV_NEE)
    r.render_samples(256)
    s = r.sums()
    r.shutdown()
    got = (s[..., :3] / s[..., 3:4]).mean(axis=(0, 1))
    h, w = sky["height"], sky["width"]
    lat = np.pi * (1.0 - (np.arange(h) + 0.5) / h)                      # scene.rs:95-111, row -> latitude of the LOOKUP direction
    d_omega = (2 * np.pi / w) * (np.pi / h) * np.sin(lat)
    cos_n = np.maximum(0.0, -np.cos(lat))                               # light arrives along -lookup (trace.rs:21); n = +y
    expect = rho / np.pi * (sky["texels"].astype(np.float64) * (cos_n * d_omega)[:, None, None]).sum(axis=(0, 1))
    _, _, weight = pb.camera_rays(cam, 2, 0)  # every sample is scaled by the cosine at the lens (camera.rs:91, worker.rs:76)
    expect = expect * weight.mean()
    assert np.all(np.abs(got - expect) <= 0.01 * expect), (got, expect)


def test_environment_nee_agrees_with_the_default_estimator_and_lowers_variance(pb, oracle_mod):
    objects = [sphere([0.0, 1.0, 0.0], 1.0, lam([0.8, 0.6, 0.4])), sphere([2.2, 0.7, 0.5], 0.7, {"type": "Mirror"}),
               sphere([0.0, -1000.0, 0.0], 1000.0, lam([0.7, 0.7, 0.7]))]
    # (1) under a sky without a sun both estimators converge to the same image.  (They cannot under a sun: the reference's
    # hemisphere sampler is not cosine-distributed, SURVEY M1, so its BSDF-sampled sky term is biased wherever the sky is not
    # uniform, while NEE integrates brdf() exactly, as the reference's own light NEE does.)
    flat = sun_sky(sun=0.08)
    desc = {"camera": CAM, "models": [], "lights": [], "sky": flat, "objects": objects}
    ref, _ = render_mean(pb, desc, 1024, seed=3)
    nee, st = render_mean(pb, desc, 1024, seed=4, extensions=pb.EXT_ENV_NEE)
    assert st["shadow_rays"] > 0
    m0, m1 = ref.mean(axis=(0, 1)), nee.mean(axis=(0, 1))
    assert np.all(np.abs(m1 - m0) <= 0.02 * m0), (m0, m1)
    # (2) under a small bright sun, sampling it directly converges much faster than waiting for BSDF samples to find it
    desc = {"camera": CAM, "models": [], "lights": [], "sky": sun_sky(), "objects": objects}
    conv_nee, _ = render_mean(pb, desc, 2048, seed=6, extensions=pb.EXT_ENV_NEE)
    conv_ref, _ = render_mean(pb, desc, 2048, seed=6)
    lo_ref, _ = render_mean(pb, desc, 16, seed=5)
    lo_nee, _ = render_mean(pb, desc, 16, seed=5, extensions=pb.EXT_ENV_NEE)
    rel = lambda x, y: float(np.sqrt(np.mean((x - y) ** 2 / (y**2 + 0.01))))  # noqa: E731
    assert rel(lo_nee, conv_nee) < 0.6 * rel(lo_ref, conv_ref)
    # (3) the default renderer (no extension) still follows the oracle on this scene
    osc = oracle_mod.Scene(desc)
    a = osc.render(CAM, 3, 0, 32)[0] / 32
    g, _ = render_mean(pb, desc, 32, seed=3)
    b = osc.render(CAM, 9, 0, 32)[0] / 32
    assert rel(g, a) <= 0.5 * rel(a, b)


def micro_desc():
    ct = {"type": "CookTorrance", "albedo": {"type": "Rgb", "rgb": [0.9, 0.7, 0.3]}, "roughness": 0.35}
    fr = {"type": "Fresnel", "refractive_index": 1.5, "diffuse": lam([0.2, 0.4, 0.8]),
          "specular": {"type": "CookTorrance", "albedo": {"type": "Rgb", "rgb": [1.0, 1.0, 1.0]}, "roughness": 0.2}}
    return {"camera": CAM, "models": [], "lights": [{"geometry": "Sphere", "v": [-3.0, 6.0, -2.0], "radius": 0.8, "colour": [1.0, 1.0, 1.0], "intensity": 15.0}],
            "sky": {"type": "Flat", "c0": [0.6, 0.7, 0.8], "c1": [0, 0, 0]},
            "objects": [sphere([-1.2, 1.0, 0.0], 1.0, ct), sphere([1.2, 1.0, 0.0], 1.0, fr), sphere([0.0, -1000.0, 0.0], 1000.0, lam([0.6, 0.6, 0.6]))]}


def test_default_ends_paths_where_the_reference_panics(pb, oracle_mod):
    # Material::sample is unimplemented for CookTorrance / Fresnel (material.rs:86): the path ends after NEE and is counted
    desc = micro_desc()
    osc = oracle_mod.Scene(desc)
    sums, counts, ctr = osc.render(CAM, 7, 0, 4)
    sc = pb.Scene(desc)
    r = pb.Renderer(CAM, sc, precision=64, seed=7)
    r.render_samples(4)
    s, st = r.sums(), r.stats()
    r.shutdown()
    assert ctr.bad_samples > 0 and abs(st["bad_samples"] - ctr.bad_samples) <= 2 + 2e-3 * ctr.bad_samples
    close = np.all(np.abs(s[..., :3] - sums) <= 1e-9 * (np.abs(sums) + 1e-3), axis=-1)
    assert close.mean() >= 0.995


@pytest.mark.parametrize("ggx", [False, True])
def test_microfacet_sampling_matches_cosine_sampling(pb, ggx):
    # two unbiased estimators of the same integrand (the reference's brdf()): importance-sampled half vectors vs a cosine lobe
    desc = micro_desc()
    extra = pb.EXT_GGX if ggx else 0
    a, sa = render_mean(pb, desc, 1024, seed=11, extensions=pb.EXT_MICROFACET | extra)
    b, sb = render_mean(pb, desc, 1024, seed=12, extensions=pb.EXT_MICROFACET_COSINE | extra)
    assert sa["bad_samples"] == 0 and sb["bad_samples"] == 0
    ma, mb = a.mean(axis=(0, 1)), b.mean(axis=(0, 1))
    assert np.all(np.abs(ma - mb) <= 0.03 * mb), (ma, mb)
    blk = lambda im: im.reshape(6, 8, 8, 8, 3).mean(axis=(1, 3))  # noqa: E731
    assert np.median(np.abs(blk(a) - blk(b)) / (blk(b) + 0.01)) < 0.03  # robust to the fireflies of the cosine estimator
    plain, _ = render_mean(pb, desc, 64, seed=11)
    assert a.mean() > plain.mean()  # paths no longer die on the two spheres


def test_all_lights_weighting(pb):
    # direct light only (max_bounces = 0): with the n-lights factor the two-light image is the sum of the one-light images
    ground = sphere([0.0, -1000.0, 0.0], 1000.0, lam([0.7, 0.7, 0.7]))
    l1 = {"geometry": "Sphere", "v": [-6.0, 5.0, 3.0], "radius": 0.5, "colour": [1.0, 0.2, 0.2], "intensity": 40.0}
    l2 = {"geometry": "Sphere", "v": [6.0, 5.0, 3.0], "radius": 0.5, "colour": [0.2, 0.2, 1.0], "intensity": 40.0}
    base = {"camera": CAM, "models": [], "sky": {"type": "Flat", "c0": [0, 0, 0], "c1": [0, 0, 0]}, "objects": [sphere([0.0, 1.0, 0.0], 1.0, lam([0.8, 0.8, 0.8])), ground]}
    both, _ = render_mean(pb, dict(base, lights=[l1, l2]), 512, seed=1, max_bounces=0, extensions=pb.EXT_NEE_ALL_LIGHTS)
    ref_both, _ = render_mean(pb, dict(base, lights=[l1, l2]), 512, seed=1, max_bounces=0)
    a, _ = render_mean(pb, dict(base, lights=[l1]), 512, seed=2, max_bounces=0)
    b, _ = render_mean(pb, dict(base, lights=[l2]), 512, seed=3, max_bounces=0)
    s = (a + b).mean(axis=(0, 1))
    assert np.all(np.abs(both.mean(axis=(0, 1)) - s) <= 0.03 * s)
    assert np.all(np.abs(2 * ref_both.mean(axis=(0, 1)) - s) <= 0.03 * s)  # the reference's estimator is low by the factor n
