#!/usr/bin/env python
"""Writes the committed fixtures under tests/golden/.

reference_kats.json  -- every known-answer vector the reference's own unit tests hold for the render path, copied by hand from
                        /root/reference (file:line given per entry).  These are the only numbers the reference pins.
oracle_vectors.npz   -- outputs of the CPU oracle (oracle/, the C++ restatement of the reference) on small seeded inputs:
                        CMJ samples, camera rays, closest hits, estimator sums, tonemapped bytes.  They are NOT reference
                        outputs (the Rust crate cannot be built here); they freeze the restatement so that
                        a later edit of the oracle cannot move the target silently, and they let the GPU tests check the CUDA
                        path without trusting the oracle build on the GPU box.

    python tools/make_golden.py        (needs oracle/liboracle.so: make -C oracle)
"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
OUT = ROOT / "tests" / "golden"

KATS = {
    "morton_code": {"cite": "src/bvh.rs:395-398", "num_bits": 4, "x": 0b1001, "y": 0b1100, "z": 0b0011,
                    "expect": (0b1100_1000_1101_0000 << 48)},
    "matrix_multiply": {"cite": "src/matrix.rs:109-135", "a": [1, 2, 3, 4, 5, 6, 7, 8, 9], "b": [9, 8, 7, 6, 5, 4, 3, 2, 1],
                        "expect": [30, 24, 18, 84, 69, 54, 138, 114, 90]},
    "switch_basis": {"cite": "src/geom.rs:352-355", "cases": [
        {"vec": [0, 1, 0], "normal": [0, 1, 0], "expect": [0, 1, 0]},
        {"vec": [0, 1, 0], "normal": [1, 0, 0], "expect": [1, 0, 0]},
        {"vec": [1, 0, 0], "normal": [0, 1, 0], "expect": [1, 0, 0]},
        {"vec": [1, 0, 0], "normal": [1, 0, 0], "expect": [0, 0, 1]}]},
    "pixel_grid_iter": {"cite": "src/worker.rs:198-234", "cases": [
        {"x0": 0, "y0": 0, "w": 3, "h": 1, "expect": [[0, 0], [1, 0], [2, 0]]},
        {"x0": 0, "y0": 0, "w": 1, "h": 3, "expect": [[0, 0], [0, 1], [0, 2]]},
        {"x0": 0, "y0": 0, "w": 2, "h": 2, "expect": [[0, 0], [1, 0], [0, 1], [1, 1]]},
        {"x0": 4, "y0": 4, "w": 2, "h": 2, "expect": [[4, 4], [5, 4], [4, 5], [5, 5]]}]},
    "sampler_properties": {"cite": "src/sampling.rs:267-308", "pattern": [2, 3], "count": 6, "seed": 42, "big_pattern": [100, 100],
                           "what": "count == m*n; same seed => same sequence; every sample inside the unit square / unit disk"},
}

SCENES = [("spheres_on_plane", 24, 16), ("bokeh_demo", 24, 16), ("teapot", 36, 24), ("bunny", 36, 24)]


def main():
    import oracle

    OUT.mkdir(parents=True, exist_ok=True)
    (OUT / "reference_kats.json").write_text(json.dumps(KATS, indent=1) + "\n")
    v = {}
    # CMJ (sampling.rs:187-264): pattern (5,5) and (1,1), two seeds
    L = oracle.lib()
    import ctypes as C

    cmj = []
    for (m, n, p) in ((5, 5, 12345), (5, 5, 0xDEADBEEF), (1, 1, 7), (2, 3, 42)):
        for s in range(m * n):
            sq, dk = np.zeros(2), np.zeros(2)
            L.orc_cmj(s, m, n, p, sq.ctypes.data_as(C.c_void_p))
            L.orc_cmj_disk(s, m, n, p, dk.ctypes.data_as(C.c_void_p))
            cmj.append([m, n, p, s, sq[0], sq[1], dk[0], dk[1]])
    v["cmj"] = np.array(cmj)
    for name, w, h in SCENES:
        osc, cam = oracle.load_scene(str(ROOT / "scenes" / f"{name}.yml"), width=w, height=h)
        o, d, wt = oracle.camera_rays(cam, 9, 0)
        v[f"{name}.rays_o"], v[f"{name}.rays_d"], v[f"{name}.rays_w"] = o, d, wt
        # fixed rays for the traversal checks are stored f32-representable, so that every instantiation sees the same ray
        o, d = o.astype(np.float32).astype(np.float64), d.astype(np.float32).astype(np.float64)
        p, t, b = osc.trace_closest(o, d, want_bary=True)
        v[f"{name}.fixed_o"], v[f"{name}.fixed_d"] = o, d
        v[f"{name}.prim"], v[f"{name}.t"], v[f"{name}.bary"] = p, t, b
        # bounce rays: leave every hit point in a seeded random direction
        rng = np.random.default_rng(3)
        hit = p != oracle.MISS
        o2 = o[hit] + d[hit] * t[hit, None]
        d2 = rng.normal(size=o2.shape)
        d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
        o2 = o2 + d2 * 1e-3
        o2, d2 = o2.astype(np.float32).astype(np.float64), d2.astype(np.float32).astype(np.float64)
        p2, t2 = osc.trace_closest(o2, d2)
        v[f"{name}.rays2_o"], v[f"{name}.rays2_d"], v[f"{name}.prim2"], v[f"{name}.t2"] = o2, d2, p2, t2
        sums, counts, _ = osc.render(cam, 5, 0, 3)
        v[f"{name}.sums"], v[f"{name}.counts"] = sums, counts
        v[f"{name}.rgb24"] = oracle.to_bytes((sums / 3.0).reshape(-1, 3)).reshape(h, w, 3)
    np.savez_compressed(OUT / "oracle_vectors.npz", **v)
    print("wrote", OUT / "reference_kats.json", "and", OUT / "oracle_vectors.npz", sum(a.nbytes for a in v.values()), "bytes raw")


if __name__ == "__main__":
    main()
