"""Seeded synthetic stand-ins for the reference's git-ignored /assets (dragon_vrip.ply, the HDRI, the
Doom PLY and the table OBJ are absent from the reference checkout).  Written where the shipped YAMLs expect them, so the
scene files are used unmodified.

    python tools/make_assets.py [dragon] [hdri] [doom] [table]

* dragon: a closed 2-manifold with the Stanford dragon's size and triangle count (871 200 vs 871 414): a tube around a (2,3)
  torus knot, displaced by seeded multi-octave noise so that it has scan-like small-scale detail.  A surface, not a random
  soup: BVH traversal cost depends on overlap, and a soup of intersecting random triangles is nothing like a scanned mesh.
* hdri: 2048x1024 Radiance .hdr, sky gradient plus a sun disk.
* doom: 300k-triangle terrain-like mesh with uchar vertex colours at the +-300 coordinates doom.yml uses (exercises
  `albedo: {type: Vertex}` and the f64 precision switch).
* table: 100k-triangle OBJ with one MTL material (exercises `material: {type: Auto}`).
"""
from __future__ import annotations

import os
import struct
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent


def write_ply(path, verts, faces, colours=None):
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    verts = np.ascontiguousarray(verts, dtype="<f4")
    faces = np.ascontiguousarray(faces, dtype="<i4")
    hdr = "ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n" % len(verts)
    if colours is not None:
        hdr += "property uchar red\nproperty uchar green\nproperty uchar blue\n"
    hdr += "element face %d\nproperty list uchar int vertex_indices\nend_header\n" % len(faces)
    with open(path, "wb") as fh:
        fh.write(hdr.encode())
        if colours is None:
            fh.write(verts.tobytes())
        else:
            rec = np.zeros(len(verts), dtype=[("p", "<f4", 3), ("c", "u1", 3)])
            rec["p"], rec["c"] = verts, colours
            fh.write(rec.tobytes())
        rec = np.zeros(len(faces), dtype=[("n", "u1"), ("i", "<i4", 3)])
        rec["n"], rec["i"] = 3, faces
        fh.write(rec.tobytes())


def _noise(u, v, rng, octaves=5):
    out = np.zeros_like(u)
    for o in range(octaves):
        f = 2 ** (o + 2)
        ph = rng.uniform(0, 2 * np.pi, size=4)
        out += (np.sin(2 * np.pi * f * u + ph[0]) * np.cos(2 * np.pi * f * v + ph[1]) + np.sin(2 * np.pi * (f + 1) * (u + v) + ph[2])
                * 0.5 * np.cos(2 * np.pi * (2 * f) * v + ph[3])) / (1.6 ** o)
    return out


def grid_faces(nu, nv, wrap_u=True, wrap_v=True):
    iu = np.arange(nu if wrap_u else nu - 1)
    iv = np.arange(nv if wrap_v else nv - 1)
    a, b = np.meshgrid(iu, iv, indexing="ij")
    a1, b1 = (a + 1) % nu, (b + 1) % nv
    i00, i10, i01, i11 = a * nv + b, a1 * nv + b, a * nv + b1, a1 * nv + b1
    f = np.stack([np.stack([i00, i10, i11], -1), np.stack([i00, i11, i01], -1)], axis=2)
    return f.reshape(-1, 3)


def make_dragon(path=ROOT / "assets/dragon_recon/dragon_vrip.ply", nu=1320, nv=330, seed=0xD7A60):
    rng = np.random.default_rng(seed)
    u, v = np.meshgrid(np.arange(nu) / nu, np.arange(nv) / nv, indexing="ij")
    t = 2 * np.pi * u

    def knot(t):
        r = 2.0 + np.cos(3 * t)
        return np.stack([r * np.cos(2 * t), np.sin(3 * t) * 1.2, r * np.sin(2 * t)], -1)

    c = knot(t)
    tan = knot(t + 1e-4) - knot(t - 1e-4)
    tan /= np.linalg.norm(tan, axis=-1, keepdims=True)
    up = np.array([0.0, 1.0, 0.0])
    n1 = np.cross(tan, up)
    n1 /= np.linalg.norm(n1, axis=-1, keepdims=True)
    n2 = np.cross(tan, n1)
    rad = 0.55 * (1.0 + 0.12 * _noise(u, v, rng))
    phi = 2 * np.pi * v
    p = c + rad[..., None] * (np.cos(phi)[..., None] * n1 + np.sin(phi)[..., None] * n2)
    # fit the Stanford dragon's bounding box (about 0.205 x 0.145 x 0.092, resting on y = 0.053)
    p -= p.min(axis=(0, 1))
    p /= p.max(axis=(0, 1))
    p = p * np.array([0.205, 0.145, 0.092]) + np.array([-0.11, 0.053, -0.045])
    write_ply(path, p.reshape(-1, 3), grid_faces(nu, nv))
    return path


def make_hdri(path=ROOT / "assets/environments/spruit_sunrise_4k.hdr", w=2048, h=1024, seed=0x5A7):
    from_top = (np.arange(h) + 0.5) / h
    sky = 0.2 + 1.8 * np.clip(1.0 - from_top * 1.6, 0.0, 1.0)
    img = np.zeros((h, w, 3), dtype=np.float32)
    img[..., 0] = (sky * 0.9)[:, None]
    img[..., 1] = (sky * 1.0)[:, None]
    img[..., 2] = (sky * 1.25)[:, None]
    img[from_top > 0.5] *= np.array([0.35, 0.3, 0.25], dtype=np.float32)  # ground half
    ys, xs = np.mgrid[0:h, 0:w]
    sun = ((xs - 0.7 * w) * (2 * np.pi / w)) ** 2 + ((ys - 0.3 * h) * (np.pi / h)) ** 2 < np.radians(1.5) ** 2
    img[sun] = np.array([5e3, 4.5e3, 4e3], dtype=np.float32)
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    mx = img.max(axis=2)
    nz = mx > 1e-32
    m, ex = np.frexp(mx)
    rgbe = np.zeros((h, w, 4), dtype=np.uint8)
    sc = np.where(nz, m * 256.0 / np.where(nz, mx, 1.0), 0.0).astype(np.float32)
    rgbe[..., :3] = np.clip(img * sc[..., None], 0, 255).astype(np.uint8)
    rgbe[..., 3] = np.where(nz, ex + 128, 0).astype(np.uint8)
    with open(path, "wb") as fh:
        fh.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n")
        fh.write(("-Y %d +X %d\n" % (h, w)).encode())
        fh.write(rgbe.tobytes())
    return path


def make_doom(path=ROOT / "assets/Doom combat scene.ply", n=388, seed=0xD00):
    rng = np.random.default_rng(seed)
    u, v = np.meshgrid(np.arange(n) / (n - 1), np.arange(n) / (n - 1), indexing="ij")
    hgt = 40.0 * _noise(u, v, rng, octaves=4)
    p = np.stack([(u - 0.5) * 600.0, hgt, (v - 0.5) * 600.0], -1)
    col = np.stack([128 + 100 * np.sin(8 * u), 128 + 100 * np.cos(6 * v), 90 + 0 * u], -1).clip(0, 255).astype(np.uint8)
    write_ply(path, p.reshape(-1, 3), grid_faces(n, n, False, False), col.reshape(-1, 3))
    return path


def make_table(path=ROOT / "assets/round_table_with_tablecloth/13493_Round_Table_with_Tablecloth_v1_L3.obj", nu=224, nv=224):
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    u, v = np.meshgrid(np.arange(nu) / nu, np.arange(nv) / (nv - 1), indexing="ij")
    r = 1.0 * np.sin(np.pi * v) ** 0.25
    p = np.stack([r * np.cos(2 * np.pi * u), 0.9 + 0.08 * np.cos(np.pi * v), r * np.sin(2 * np.pi * u)], -1).reshape(-1, 3)
    f = grid_faces(nu, nv, True, False) + 1
    (path.parent / "table.mtl").write_text("newmtl wood\nKd 0.55 0.35 0.2\n")
    with open(path, "w") as fh:
        fh.write("mtllib table.mtl\no table\nusemtl wood\n")
        np.savetxt(fh, p, fmt="v %.6f %.6f %.6f")
        np.savetxt(fh, f, fmt="f %d %d %d")
    return path


def make_vertex_colours(path=ROOT / "tests/golden/vertex_colours.ply", n=28, seed=0xC0105):
    """A small vertex-coloured terrain inside +-8 units (committed fixture): the f32 production path's per-vertex albedo and smooth
    normals against the oracle (doom.yml, the only reference scene with `albedo: {type: Vertex}`, works at +-700 and resolves to the
    mixed-precision instantiation, so it never reaches the f32 mirrors)."""
    rng = np.random.default_rng(seed)
    u, v = np.meshgrid(np.arange(n) / (n - 1), np.arange(n) / (n - 1), indexing="ij")
    hgt = 0.6 * _noise(u, v, rng, octaves=3)
    p = np.stack([(u - 0.5) * 12.0, hgt, (v - 0.5) * 12.0], -1)
    col = np.stack([128 + 120 * np.sin(9 * u), 128 + 120 * np.cos(7 * v), 128 + 120 * np.sin(5 * (u + v))], -1).clip(0, 255).astype(np.uint8)
    write_ply(path, p.reshape(-1, 3), grid_faces(n, n, False, False), col.reshape(-1, 3))
    return path


MAKERS = {"dragon": make_dragon, "hdri": make_hdri, "doom": make_doom, "table": make_table, "vertex_colours": make_vertex_colours}


def ensure(*names):
    os.chdir(ROOT)
    out = {}
    for n in names:
        fn = MAKERS[n]
        default = fn.__defaults__[0]
        if not Path(default).exists():
            # written under a private name and renamed: several ranks of one bench may get here at the same time
            tmp = Path(default).with_name(Path(default).name + f".tmp{os.getpid()}")
            fn(tmp)
            os.replace(tmp, default)
        out[n] = str(default)
    return out


if __name__ == "__main__":
    which = sys.argv[1:] or list(MAKERS)
    for n in which:
        print(n, MAKERS[n]())
