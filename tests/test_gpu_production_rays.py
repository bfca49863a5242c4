"""Fixed-ray parity of the PRODUCTION traversal kernels (run on the B200 box with -m gpu).

north_star's first correctness bar -- closest-hit primitive id exact, t within 1e-4 relative -- is checked here on the kernels a
render really launches: `paths_trace_closest(precision=32)` stages the rays into a path-state buffer and runs `k_extend_refill`
(two-wide nodes, scenes up to 64k primitives) or `k_extend4s` (four-wide nodes, parked leaves) with the grid a render uses;
`PATHS_TRACE_CAMERA` is the routine of the fused camera kernel; `paths_trace_occluded` runs `k_shadow_refill` / `k_shadow4` over
staged visibility records and reads the occlusion flag back from the radiance record the kernel adds to.

Ray sets per scene, at BASELINE's resolution: the first camera sample of all 1920x1080 pixels (2 073 600 rays) plus 1 000 000 incoherent
rays leaving the visible surfaces in random directions (what bounce and shadow rays look like).  The oracle traces the rays the f32
kernels see (rounded to f32).  Every id difference must be explained by the oracle itself: the two candidates tie within 1e-4 relative,
or the ray passes within f32 position error of a triangle edge (|barycentric| < 2e-3 in the oracle's own f64 arithmetic).
Reference: /root/reference/src/bvh.rs:78-140 (closest hit), src/trace.rs:61-66 (shadow query).
"""
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
W, H = 1920, 1080
N_INCOHERENT = 1_000_000


@pytest.fixture(scope="module")
def pb():
    import paths_b200

    if paths_b200.device_count() < 1:
        pytest.fail("no CUDA device: the product has no CPU path")
    sys.path.insert(0, str(ROOT / "tools"))
    import make_assets

    make_assets.ensure("dragon", "hdri", "table")
    return paths_b200


class Case:
    """One scene at 1080p: oracle scene, ray sets (already rounded to f32) and the oracle's answers, computed once per module."""

    def __init__(self, oracle_mod, name):
        self.name = name
        self.osc, self.cam = oracle_mod.load_scene(f"scenes/{name}.yml", width=W, height=H)
        o, d, _ = oracle_mod.camera_rays(self.cam, 11, 0)
        o, d = o.astype(np.float32).astype(np.float64), d.astype(np.float32).astype(np.float64)
        p, t, b = self.osc.trace_closest(o, d, want_bary=True)
        self.camera = (o, d, p, t, b)
        rng = np.random.default_rng(17)
        hit = np.nonzero(p != oracle_mod.MISS)[0]
        sel = rng.choice(hit, size=min(N_INCOHERENT, len(hit)), replace=False)
        d2 = rng.normal(size=(len(sel), 3))
        d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
        o2 = o[sel] + d[sel] * t[sel, None] + d2 * 1e-3
        o2, d2 = o2.astype(np.float32).astype(np.float64), d2.astype(np.float32).astype(np.float64)
        p2, t2, b2 = self.osc.trace_closest(o2, d2, want_bary=True)
        self.incoherent = (o2, d2, p2, t2, b2)


@pytest.fixture(scope="module")
def cases(oracle_mod):
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = Case(oracle_mod, name)
        return cache[name]

    return get


def tri_bary(geom, o, d):
    """The reference's triangle test (src/geom.rs:265-303) in numpy f64 -> (t, min barycentric), whatever their signs."""
    a, b, c, n = geom[0:3], geom[3:6], geom[6:9], geom[9:12]
    cos_theta = n @ d
    if cos_theta == 0.0:
        return np.inf, -np.inf
    t = (n @ a - n @ o) / cos_theta
    p = o + d * t
    area = n @ np.cross(b - a, c - a)
    x = n @ np.cross(b - p, c - p) / area
    y = n @ np.cross(c - p, a - p) / area
    return t, min(x, y, 1.0 - x - y)


def unresolvable(geom, o, d, tol_edge=2e-3):
    """True when f32 rays cannot decide whether this triangle is hit: the ray passes an edge at less than 16 ulps of the coordinates
    (measured perpendicular to the ray: |barycentric| x edge length x cos of incidence -- a grazing ray moves a long way inside the
    triangle's plane for one ulp of its origin), or within 2e-3 barycentric units, or the plane is crossed within 16 ulps of the origin."""
    t, mb = tri_bary(geom, o, d)
    a, b, c, n = geom[0:3], geom[3:6], geom[6:9], geom[9:12]
    ulp = 2.0**-23 * max(np.abs(o).max(), 1.0)
    if abs(t) <= 16 * ulp:
        return True
    if abs(mb) < tol_edge:
        return True
    edge = min(np.linalg.norm(b - a), np.linalg.norm(c - a), np.linalg.norm(c - b))
    cos = abs(n @ d) / (np.linalg.norm(n) * np.linalg.norm(d))
    return abs(mb) * edge * cos <= 16 * ulp


def explain_id_differences(oracle_mod, osc, gsc, o, d, p0, b0, p1):
    """Returns the indices of differing ids that are neither ties within 1e-4 relative nor cases f32 rays cannot resolve."""
    unexplained = []
    for i in np.nonzero(p0 != p1)[0]:
        ta = osc.intersect_prim(p0[i], o[i], d[i]) if p0[i] != oracle_mod.MISS else None
        tb = osc.intersect_prim(p1[i], o[i], d[i]) if p1[i] != oracle_mod.MISS else None
        if ta is not None and tb is not None and abs(ta - tb) <= 1e-4 * abs(ta):
            continue  # tie: two primitives at the same distance (shared edge or coincident faces); either is a closest hit
        ok = False
        for p in (p0[i], p1[i]):  # the oracle's hit, or the GPU's, is a graze / a crossing at the origin
            if p != oracle_mod.MISS and not ok:
                is_tri, geom = gsc.prim_geometry(p)
                ok = is_tri and unresolvable(geom, o[i], d[i])
        if not ok:
            unexplained.append(int(i))
    return unexplained


def check_closest(pb, oracle_mod, case, gsc, rays, kernel, max_diff_frac):
    o, d, p0, t0, b0 = rays
    p1, t1 = gsc.find_intersection(o, d, precision=kernel)
    same = p0 == p1
    hit = same & (p0 != oracle_mod.MISS)
    assert hit.any()
    # t within 1e-4 relative (north_star) -- a forward bound that f32 rays cannot meet everywhere: (a) a ray that ends a few
    # millimetres from its origin (surface rays hitting the neighbouring triangle) knows its own origin only to an ulp of the scene's
    # coordinates, so its distance is held to 16 of those ulps instead; (b) a ray nearly parallel to the primitive it hits is
    # ill-conditioned in t, and is held to the backward error: the reported hit point lies ON the oracle's primitive within 16 ulps.
    # Both classes are rare and counted; everything else must meet the relative bound.
    idx = np.nonzero(hit)[0]
    err = np.abs(t1[idx] - t0[idx])
    fwd = err <= 1e-4 * np.abs(t0[idx])
    coherent = rays is case.camera
    tier1 = float(fwd.mean())
    if coherent:
        assert tier1 > 0.9999, f"{case.name} kernel {kernel}: only {tier1:.6f} of the camera-ray distances within 1e-4 relative"
    # incoherent surface rays: how many end within millimetres of their origin depends on the mesh (teapot 0.9996 within the relative
    # bound, the 871k sub-pixel triangles of the dragon stand-in 0.998, environment.yml 0.98); those are held to the origin's resolution
    fwd |= err <= 16 * 2.0**-23 * np.maximum(np.abs(o[idx]).max(axis=1), 1.0)
    tier2 = float(fwd.mean())
    print(f"[distances] {case.name} kernel {kernel} {'camera' if coherent else 'incoherent'}: {tier1:.6f} within 1e-4 relative, "
          f"{tier2:.6f} within that or 16 ulps of the origin, {int((~fwd).sum())} checked for backward error")
    assert tier2 > 0.999, f"{case.name} kernel {kernel}: only {tier2:.6f} of the distances within 1e-4 relative or 16 ulps of the origin"
    for i in idx[~fwd]:
        is_tri, geom = gsc.prim_geometry(p0[i])
        point = o[i] + d[i] * t1[i]
        ulp = 2.0**-23 * max(np.abs(o[i]).max(), np.abs(point).max(), 1.0)
        if is_tri:
            nrm = geom[9:12] / np.linalg.norm(geom[9:12])
            off = abs(nrm @ (point - geom[0:3]))
        else:
            off = abs(np.linalg.norm(point - geom[0:3]) - geom[3])
            ulp = max(ulp, 2.0**-52 * geom[3] * 4)  # spheres are intersected in f64: the 1e6-radius "planes" resolve to ~1e-9
        assert off <= 16 * ulp, (f"{case.name} kernel {kernel}: ray {i} t0 {t0[i]!r} t1 {t1[i]!r} prim {p0[i]}: the GPU's hit point is {off:.3g} "
                                 f"off the primitive ({off / ulp:.1f} ulp)")
    assert np.all(np.isinf(t1[same & ~hit]))
    n_diff = int((~same).sum())
    assert n_diff <= max_diff_frac * len(p0), f"{case.name}: {n_diff} of {len(p0)} ids differ"
    bad = explain_id_differences(oracle_mod, case.osc, gsc, o, d, p0, b0, p1)
    assert not bad, f"{case.name} kernel {kernel}: {len(bad)} of {n_diff} id differences are neither ties nor edge grazes, e.g. ray {bad[:5]}"
    return p1, t1


SMALL = ["teapot", "bunny"]          # two-wide kernels in a render (<= 64k primitives)
LARGE = ["dragon", "environment"]    # four-wide kernels in a render


def assert_no_stack_overflow(pb, gsc, cam):
    for prec in (32, 48):
        r = pb.Renderer(cam, gsc, precision=prec, samples_per_pass=1)
        st = r.stats()
        r.shutdown()
        assert st["stack_overflows"] == 0


@pytest.mark.parametrize("name", SMALL + LARGE)
@pytest.mark.parametrize("builder", ["host_sah", "device_lbvh"])
def test_production_closest_hit_kernels_on_fixed_rays(pb, oracle_mod, cases, name, builder):
    case = cases(name)
    gsc = pb.Scene.load(f"scenes/{name}.yml", width=W, height=H)
    if builder == "device_lbvh":
        gsc.build_bvh(device=True)
    # sub-pixel triangles of the 871k-triangle stand-in mesh: ~1e-3 of the rays pass within f32 error of an edge
    frac = 5e-3 if name in LARGE else 2e-4
    # the kernel a render launches for bounce rays, the camera-segment routine, and the other tree width for good measure
    kernels = [pb.PRECISION_F32, pb.TRACE_CAMERA, pb.TRACE_FOUR_WIDE if name in SMALL else pb.TRACE_TWO_WIDE]
    for rays in (case.camera, case.incoherent):
        got = [check_closest(pb, oracle_mod, case, gsc, rays, kernel, frac) for kernel in kernels]
        # exact ties go to the lower primitive id whatever order a kernel visits the tree in (dev_trace.cuh `closer`): the plain-batch,
        # lane-refill two-wide and parked-leaf four-wide kernels must agree bit for bit, on ids AND distances
        for p1, t1 in got[1:]:
            assert np.array_equal(p1, got[0][0]) and np.array_equal(t1, got[0][1]), f"{name}: production kernels disagree with each other"
    r = pb.Renderer(case.cam, gsc, precision=32, samples_per_pass=1)
    st = r.stats()
    r.shutdown()
    assert st["extend_kernel"] == (3 if name in SMALL else 16) and st["camera_kernel"] == 1 and st["stack_overflows"] == 0
    assert_no_stack_overflow(pb, gsc, case.cam)
    if builder == "device_lbvh":
        assert gsc.info["bvh_builder"] == 1 and gsc.check_bvh() == 0


def light_rays(case, oracle_mod, rng, light):
    """Visibility rays as trace_ray makes them (trace.rs:53-66): from surface points toward points on the light's sphere."""
    o2, _, _, _, _ = case.incoherent
    centre, radius = np.asarray(light["v"], dtype=np.float64), float(light.get("radius", 0.0))
    n = rng.normal(size=o2.shape)
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    target = centre + n * radius
    d = target - o2
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o2, d.astype(np.float32).astype(np.float64)


def prim_tables(osc, prims):
    """is_light, entity of the given primitive ids (oracle's flattened item list, scene.rs:143-170)."""
    uniq = np.unique(prims)
    il, en = {}, {}
    for p in uniq:
        pi = osc.prim_info(int(p))
        il[int(p)], en[int(p)] = bool(pi.is_light), int(pi.entity)
    return np.array([il[int(p)] for p in prims], dtype=bool), np.array([en[int(p)] for p in prims], dtype=np.int64)


@pytest.mark.parametrize("name", SMALL + LARGE)
def test_production_visibility_kernels_on_fixed_rays(pb, oracle_mod, cases, name):
    import yaml

    case = cases(name)
    gsc = pb.Scene.load(f"scenes/{name}.yml", width=W, height=H)
    lights = yaml.safe_load(open(ROOT / "scenes" / f"{name}.yml"))["lights"]
    rng = np.random.default_rng(23)
    if lights:
        assert lights[0]["geometry"]["type"] == "Sphere"
        c = lights[0]["geometry"]["center"]
        lt = {"v": [c["x"], c["y"], c["z"]], "radius": lights[0]["geometry"]["radius"]}
        o, d = light_rays(case, oracle_mod, rng, lt)
        light = 0
    else:  # no light in the scene: environment-style records, every hit occludes
        o, d = case.incoherent[0], case.incoherent[1]
        light = None
    p0, t0, b0 = case.osc.trace_closest(o, d, want_bary=True)
    hit = p0 != oracle_mod.MISS
    exp = hit.copy()
    if light is not None:
        il, en = prim_tables(case.osc, p0[hit])
        own = np.zeros(len(p0), dtype=bool)
        own[np.nonzero(hit)[0]] = il & (en == light)
        exp &= ~own  # trace.rs:61-66: occluded <=> the closest hit belongs to another entity than the light
        assert own.mean() > 0.02, "the ray set never reaches the light"
    assert 0.05 < exp.mean() < 0.999
    # the parity instantiation runs the reference's rule on the reference's tree: flags equal the oracle's, no exceptions
    occ64 = gsc.occluded(o[:200000], d[:200000], light=light, precision=64)
    assert np.array_equal(occ64.astype(bool), exp[:200000])
    frac = 5e-3 if name in LARGE else 2e-4
    for kernel in (pb.PRECISION_F32, pb.TRACE_FOUR_WIDE if name in SMALL else pb.TRACE_TWO_WIDE, pb.TRACE_PLAIN, pb.PRECISION_MIXED):
        occ = gsc.occluded(o, d, light=light, precision=kernel)
        assert set(np.unique(occ)) <= {0, 1}, "a radiance record was not written exactly once"
        diff = np.nonzero(occ.astype(bool) != exp)[0]
        assert len(diff) <= frac * len(exp), f"{name} kernel {kernel}: {len(diff)} of {len(exp)} flags differ"
        if kernel == pb.PRECISION_MIXED:
            # f64 primitive tests over the production tree: the predicate can only differ where an occluder ties with the light's surface
            assert len(diff) <= 1e-5 * len(exp) + 2, f"{name} mixed: {len(diff)} flags differ"
            continue
        # production f32: each difference must be explained by the oracle -- the deciding triangle is grazed within f32 position
        # error of an edge, or the occluder sits within 1e-4 relative of the ray's start or of the light's surface
        q, tq = gsc.find_intersection(o[diff], d[diff], precision=pb.PRECISION_F32)
        unexplained = []
        for j, i in enumerate(diff):
            ok = False
            for prim in {int(p0[i]), int(q[j])} - {int(oracle_mod.MISS)}:
                is_tri, geom = gsc.prim_geometry(prim)
                if is_tri:
                    tt, _ = tri_bary(geom, o[i], d[i])
                    ok = ok or unresolvable(geom, o[i], d[i]) or abs(tt) < 2e-4
            if hit[i] and t0[i] < 2e-4:
                ok = True  # the occluder is the surface the ray starts on (origin offset is 1e-3 along a random direction)
            if not ok:
                unexplained.append(int(i))
        assert not unexplained, f"{name} kernel {kernel}: {len(unexplained)} of {len(diff)} flag differences unexplained, e.g. {unexplained[:5]}"
    assert_no_stack_overflow(pb, gsc, case.cam)
