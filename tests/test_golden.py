"""The committed fixtures of tests/golden/ (made by tools/make_golden.py).

CPU part (-m "not gpu"): the oracle still reproduces the reference's own known-answer vectors (reference_kats.json) and its own
frozen outputs (oracle_vectors.npz).  GPU part: the CUDA path, through the C ABI, reproduces the frozen vectors directly -- bit
for bit in the f64 instantiation, within the production bars in f32 -- without running the oracle at all."""
import ctypes as C
import json
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).resolve().parent / "golden"
SCENES = [("spheres_on_plane", 24, 16), ("bokeh_demo", 24, 16), ("teapot", 36, 24), ("bunny", 36, 24)]


@pytest.fixture(scope="module")
def kats():
    return json.loads((GOLD / "reference_kats.json").read_text())


@pytest.fixture(scope="module")
def vec():
    return np.load(GOLD / "oracle_vectors.npz")


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_reference_kats(oracle_mod, kats):
    L = oracle_mod.lib()
    k = kats["morton_code"]
    assert L.orc_morton_code(k["num_bits"], k["x"], k["y"], k["z"]) == k["expect"]
    k = kats["matrix_multiply"]
    out = np.zeros(9)
    L.orc_matmul3(_p(np.array(k["a"], dtype=np.float64)), _p(np.array(k["b"], dtype=np.float64)), _p(out))
    assert out.tolist() == [float(x) for x in k["expect"]]
    for c in kats["switch_basis"]["cases"]:
        out = np.zeros(3)
        L.orc_switch_basis(_p(np.array(c["vec"], dtype=np.float64)), _p(np.array(c["normal"], dtype=np.float64)), _p(out))
        assert out.tolist() == [float(x) for x in c["expect"]]
    for c in kats["pixel_grid_iter"]["cases"]:
        out = np.zeros(c["w"] * c["h"] * 2, dtype=np.uint32)
        L.orc_pixel_grid(c["x0"], c["y0"], c["w"], c["h"], _p(out))
        assert out.reshape(-1, 2).tolist() == c["expect"]
    k = kats["sampler_properties"]
    m, n = k["pattern"]
    seq = []
    for s in range(m * n):
        sq = np.zeros(2)
        L.orc_cmj(s, m, n, k["seed"], _p(sq))
        seq.append(tuple(sq))
    assert len(seq) == k["count"] and len(set(seq)) == k["count"]
    m, n = k["big_pattern"]
    for s in range(0, m * n, 37):
        sq, dk = np.zeros(2), np.zeros(2)
        L.orc_cmj(s, m, n, k["seed"], _p(sq))
        L.orc_cmj_disk(s, m, n, k["seed"], _p(dk))
        assert 0.0 <= sq[0] <= 1.0 and 0.0 <= sq[1] <= 1.0 and dk[0] ** 2 + dk[1] ** 2 <= 1.0


def test_oracle_reproduces_its_frozen_vectors(oracle_mod, vec):
    L = oracle_mod.lib()
    for row in vec["cmj"]:
        m, n, p, s = (int(x) for x in row[:4])
        sq, dk = np.zeros(2), np.zeros(2)
        L.orc_cmj(s, m, n, p, _p(sq))
        L.orc_cmj_disk(s, m, n, p, _p(dk))
        assert sq.tolist() == row[4:6].tolist() and dk.tolist() == row[6:8].tolist()
    for name, w, h in SCENES:
        osc, cam = oracle_mod.load_scene(f"scenes/{name}.yml", width=w, height=h)
        o, d, wt = oracle_mod.camera_rays(cam, 9, 0)
        assert np.array_equal(o, vec[f"{name}.rays_o"]) and np.array_equal(d, vec[f"{name}.rays_d"]) and np.array_equal(wt, vec[f"{name}.rays_w"])
        p, t, b = osc.trace_closest(vec[f"{name}.fixed_o"], vec[f"{name}.fixed_d"], want_bary=True)
        assert np.array_equal(p, vec[f"{name}.prim"]) and np.array_equal(t, vec[f"{name}.t"]) and np.array_equal(b, vec[f"{name}.bary"])
        p2, t2 = osc.trace_closest(vec[f"{name}.rays2_o"], vec[f"{name}.rays2_d"])
        assert np.array_equal(p2, vec[f"{name}.prim2"]) and np.array_equal(t2, vec[f"{name}.t2"])
        sums, counts, _ = osc.render(cam, 5, 0, 3)
        # libm (sin, cos, pow, acos) may differ by an ulp between machines: sums are compared to 1e-12, not bit for bit
        assert np.allclose(sums, vec[f"{name}.sums"], rtol=1e-12, atol=1e-15) and np.array_equal(counts, vec[f"{name}.counts"])


@pytest.mark.gpu
@pytest.mark.parametrize("name,w,h", SCENES)
def test_cuda_path_reproduces_the_frozen_vectors(name, w, h, vec):
    import paths_b200 as pb

    if pb.device_count() < 1:
        pytest.fail("no CUDA device: the product has no CPU path")
    sc = pb.Scene.load(f"scenes/{name}.yml", width=w, height=h)
    o, d, wt = pb.camera_rays(sc.camera, 9, 0, precision=64)
    assert np.allclose(o, vec[f"{name}.rays_o"], rtol=0, atol=1e-12) and np.allclose(d, vec[f"{name}.rays_d"], rtol=0, atol=1e-12)
    miss = 0xFFFFFFFF
    for rays_o, rays_d, kp, kt in ((f"{name}.fixed_o", f"{name}.fixed_d", f"{name}.prim", f"{name}.t"),
                                   (f"{name}.rays2_o", f"{name}.rays2_d", f"{name}.prim2", f"{name}.t2")):
        ro, rd, p0, t0 = vec[rays_o], vec[rays_d], vec[kp], vec[kt]
        hit = p0 != miss
        p, t = sc.find_intersection(ro, rd, precision=64)        # the reference's tree and order: everything bit-equal
        assert np.array_equal(p, p0) and np.array_equal(t[hit], t0[hit])
        p, t = sc.find_intersection(ro, rd, precision=pb.PRECISION_MIXED)  # production tree, f64 primitives: distances bit-equal
        assert np.array_equal(p != miss, hit) and np.array_equal(t[hit], t0[hit]) and (p == p0).mean() > 0.999
        p, t = sc.find_intersection(ro, rd, precision=32)        # production: ids equal away from ties, t within 1e-4 relative
        same = p == p0
        assert same.mean() > 0.995
        assert np.all(np.abs(t[same & hit] - t0[same & hit]) <= 1e-4 * np.abs(t0[same & hit]))
    r = pb.Renderer(sc.camera, sc, precision=64, seed=5)
    r.render_samples(3)
    s = r.sums()
    r.shutdown()
    g = vec[f"{name}.sums"]
    close = np.all(np.abs(s[..., :3] - g) <= 1e-9 * np.abs(g) + 1e-12, axis=-1)
    assert close.mean() >= 0.99 and np.array_equal(s[..., 3], vec[f"{name}.counts"])
