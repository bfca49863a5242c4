"""Scene-file readers for the CPU oracle (TEST INFRASTRUCTURE ONLY).

Python restatement of the reference's scene I/O: the serde schema of
``src/serde.rs:65-385`` (via PyYAML), the OBJ reader ``src/obj.rs:8-67`` (tobj 2.0.0,
``Cargo.lock``), the PLY reader ``src/ply.rs:11-78`` (ply-rs 0.1.2) and the Radiance HDR
decode of ``src/serde.rs:367-382`` (image 0.23.4 ``hdr::HdrDecoder``).  The three crates
are third-party and absent from /root/reference; their published behaviour is restated
here (f32 parsing of coordinates, fan triangulation, RGBE -> f32 as ``m * 2^(e-136)``).

The product has its own C++ readers (paths_b200/csrc/host); tests compare the two.
Nothing under paths_b200/ imports this module.
"""
from __future__ import annotations

import os
import struct
from pathlib import Path

import numpy as np
import yaml

# ----------------------------------------------------------------------------- OBJ (src/obj.rs, tobj 2.0.0)


def _parse_mtl(path):
    """newmtl / Kd only: src/obj.rs:26-29 keeps nothing but the diffuse colour."""
    mats, order, cur = {}, [], None
    try:
        with open(path, "r", errors="replace") as fh:
            for line in fh:
                tok = line.split()
                if not tok or tok[0].startswith("#"):
                    continue
                if tok[0] == "newmtl":
                    cur = " ".join(tok[1:])
                    mats[cur] = [0.0, 0.0, 0.0]
                    order.append(cur)
                elif tok[0] == "Kd" and cur is not None and len(tok) >= 4:
                    mats[cur] = [float(np.float32(tok[1])), float(np.float32(tok[2])), float(np.float32(tok[3]))]
    except OSError:
        pass
    return mats, order


def load_obj(path):
    """Returns a list of models (one per OBJ object/group that has faces), like tobj::load_obj(file, true)."""
    positions = []
    models = []
    cur_faces = []
    cur_mat = None
    mats = {}

    def flush():
        nonlocal cur_faces
        if cur_faces:
            faces = np.asarray(cur_faces, dtype=np.int64)
            # tobj re-indexes the vertices used by each mesh in order of first use
            flat = faces.reshape(-1)
            uniq, first = np.unique(flat, return_index=True)
            order = uniq[np.argsort(first)]
            remap = np.full(len(positions), -1, dtype=np.int64)
            remap[order] = np.arange(len(order))
            verts = np.asarray(positions, dtype=np.float32)[order].astype(np.float64)
            material = None
            if cur_mat is not None and cur_mat in mats:
                material = {"type": "Lambertian", "albedo": {"type": "Rgb", "rgb": mats[cur_mat]}}
            models.append(
                {
                    "vertices": np.ascontiguousarray(verts),
                    "faces": np.ascontiguousarray(remap[faces].astype(np.uint32)),
                    "vertex_colours": None,
                    "material": material,
                }
            )
        cur_faces = []

    base = os.path.dirname(path)
    with open(path, "r", errors="replace") as fh:
        for line in fh:
            tok = line.split()
            if not tok or tok[0].startswith("#"):
                continue
            k = tok[0]
            if k == "v":
                positions.append((np.float32(tok[1]), np.float32(tok[2]), np.float32(tok[3])))
            elif k == "f":
                idx = []
                for t in tok[1:]:
                    vi = int(t.split("/")[0])
                    idx.append(vi - 1 if vi > 0 else len(positions) + vi)
                for i in range(1, len(idx) - 1):  # fan triangulation (triangulate = true)
                    cur_faces.append((idx[0], idx[i], idx[i + 1]))
            elif k in ("o", "g"):
                flush()
            elif k == "usemtl":
                name = " ".join(tok[1:])
                if cur_faces and name != cur_mat:
                    flush()
                cur_mat = name
            elif k == "mtllib":
                m, _ = _parse_mtl(os.path.join(base, " ".join(tok[1:])))
                mats.update(m)
    flush()
    return models


# ----------------------------------------------------------------------------- PLY (src/ply.rs, ply-rs 0.1.2)

_PLY_TYPES = {
    "char": "b", "int8": "b", "uchar": "B", "uint8": "B", "short": "h", "int16": "h", "ushort": "H", "uint16": "H",
    "int": "i", "int32": "i", "uint": "I", "uint32": "I", "float": "f", "float32": "f", "double": "d", "float64": "d",
}


def load_ply(path):
    with open(path, "rb") as fh:
        data = fh.read()
    end = data.index(b"end_header")
    end = data.index(b"\n", end) + 1
    header = data[:end].decode("ascii", "replace").splitlines()
    fmt = None
    elements = []  # (name, count, [(kind, name, type...)])
    for line in header:
        tok = line.split()
        if not tok:
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            elements.append([tok[1], int(tok[2]), []])
        elif tok[0] == "property":
            if tok[1] == "list":
                elements[-1][2].append(("list", tok[4], tok[2], tok[3]))
            else:
                elements[-1][2].append(("scalar", tok[2], tok[1]))
    body = data[end:]
    vertices = colours = None
    faces = []
    if fmt == "ascii":
        tokens = body.split()
        pos = 0
        for name, count, props in elements:
            rows = []
            for _ in range(count):
                row = {}
                for p in props:
                    if p[0] == "scalar":
                        row[p[1]] = tokens[pos]
                        pos += 1
                    else:
                        n = int(tokens[pos])
                        pos += 1
                        row[p[1]] = [int(t) for t in tokens[pos:pos + n]]
                        pos += n
                rows.append(row)
            if name == "vertex":
                vertices = np.array([[np.float32(r["x"]), np.float32(r["y"]), np.float32(r["z"])] for r in rows], dtype=np.float32)
                if rows and all(c in rows[0] for c in ("red", "green", "blue")):
                    colours = np.array([[int(r["red"]), int(r["green"]), int(r["blue"])] for r in rows], dtype=np.float64) / 255.0
            elif name == "face":
                key = "vertex_indices" if (rows and "vertex_indices" in rows[0]) else "vertex_index"
                faces = [r[key][:3] for r in rows]
    else:
        endian = "<" if fmt == "binary_little_endian" else ">"
        pos = 0
        for name, count, props in elements:
            if all(p[0] == "scalar" for p in props):
                dt = np.dtype([(p[1], endian + _PLY_TYPES[p[2]]) for p in props])
                arr = np.frombuffer(body, dtype=dt, count=count, offset=pos)
                pos += dt.itemsize * count
                if name == "vertex":
                    vertices = np.stack([arr["x"], arr["y"], arr["z"]], axis=1).astype(np.float32)
                    if all(c in arr.dtype.names for c in ("red", "green", "blue")):
                        colours = np.stack([arr["red"], arr["green"], arr["blue"]], axis=1).astype(np.float64) / 255.0
            else:
                # fast path: one list property with a constant length of 3 (triangles)
                if len(props) == 1 and count > 0:
                    cnt_t, idx_t = _PLY_TYPES[props[0][2]], _PLY_TYPES[props[0][3]]
                    rec = np.dtype([("n", endian + cnt_t), ("i", endian + idx_t, (3,))])
                    if pos + rec.itemsize * count <= len(body):
                        arr = np.frombuffer(body, dtype=rec, count=count, offset=pos)
                        if np.all(arr["n"] == 3):
                            pos += rec.itemsize * count
                            if name == "face":
                                faces = arr["i"].astype(np.int64)
                            continue
                rows = []
                for _ in range(count):
                    row = {}
                    for p in props:
                        if p[0] == "scalar":
                            c = _PLY_TYPES[p[2]]
                            row[p[1]] = struct.unpack_from(endian + c, body, pos)[0]
                            pos += struct.calcsize(c)
                        else:
                            c, it = _PLY_TYPES[p[2]], _PLY_TYPES[p[3]]
                            n = struct.unpack_from(endian + c, body, pos)[0]
                            pos += struct.calcsize(c)
                            row[p[1]] = list(struct.unpack_from(endian + str(n) + it, body, pos))
                            pos += struct.calcsize(it) * n
                    rows.append(row)
                if name == "face":
                    key = "vertex_indices" if (rows and "vertex_indices" in rows[0]) else "vertex_index"
                    faces = [r[key][:3] for r in rows]
    return {
        "vertices": np.ascontiguousarray(vertices.astype(np.float64)),
        "faces": np.ascontiguousarray(np.asarray(faces, dtype=np.uint32).reshape(-1, 3)),
        "vertex_colours": None if colours is None else np.ascontiguousarray(colours),
        "material": None,
    }


# ----------------------------------------------------------------------------- Radiance HDR (image 0.23.4 hdr::HdrDecoder)


def load_hdr(path):
    """Returns (width, height, float32 texels[h, w, 3]).  RGBE -> f32 as m * 2^(e - 136); e == 0 -> 0."""
    with open(path, "rb") as fh:
        data = fh.read()
    pos = 0
    width = height = None
    while True:
        nl = data.index(b"\n", pos)
        line = data[pos:nl].decode("ascii", "replace").strip()
        pos = nl + 1
        tok = line.split()
        if len(tok) == 4 and tok[0] in ("-Y", "+Y") and tok[2] in ("+X", "-X"):
            height, width = int(tok[1]), int(tok[3])
            break
    rgbe = np.zeros((height, width, 4), dtype=np.uint8)
    for y in range(height):
        if 8 <= width <= 0x7FFF and data[pos] == 2 and data[pos + 1] == 2 and (data[pos + 2] & 0x80) == 0:
            pos += 4
            for c in range(4):
                x = 0
                while x < width:
                    n = data[pos]
                    pos += 1
                    if n > 128:
                        n -= 128
                        rgbe[y, x:x + n, c] = data[pos]
                        pos += 1
                    else:
                        rgbe[y, x:x + n, c] = np.frombuffer(data, dtype=np.uint8, count=n, offset=pos)
                        pos += n
                    x += n
        else:
            rgbe[y] = np.frombuffer(data, dtype=np.uint8, count=width * 4, offset=pos).reshape(width, 4)
            pos += width * 4
    e = rgbe[..., 3].astype(np.int32)
    scale = np.where(e == 0, np.float32(0.0), np.exp2((e - 136).astype(np.float32))).astype(np.float32)
    texels = rgbe[..., :3].astype(np.float32) * scale[..., None]
    return width, height, np.ascontiguousarray(texels.astype(np.float32))


def write_hdr(path, texels):
    """Flat (non-RLE) RGBE writer used by the synthetic-asset generator."""
    t = np.asarray(texels, dtype=np.float32)
    h, w, _ = t.shape
    mx = t.max(axis=2)
    e = np.zeros((h, w), dtype=np.int32)
    nz = mx > 1e-32
    m, ex = np.frexp(mx[nz])
    rgbe = np.zeros((h, w, 4), dtype=np.uint8)
    sc = np.zeros((h, w), dtype=np.float32)
    sc[nz] = (m * 256.0 / mx[nz]).astype(np.float32)
    e[nz] = ex + 128
    rgbe[..., :3] = np.clip(t * sc[..., None], 0, 255).astype(np.uint8)
    rgbe[..., 3] = np.where(nz, e, 0).astype(np.uint8)
    with open(path, "wb") as fh:
        fh.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n")
        fh.write(("-Y %d +X %d\n" % (h, w)).encode())
        fh.write(rgbe.tobytes())


# ----------------------------------------------------------------------------- YAML schema (src/serde.rs)


def _vec(d):
    return [float(d["x"]), float(d["y"]), float(d["z"])]


def _col(d):
    return [float(d["r"]), float(d["g"]), float(d["b"])]


def _albedo(d):
    # src/serde.rs:42-47 is internally tagged; the stale shipped scenes omit the tag (superset: untagged => Rgb)
    t = d.get("type", "Rgb")
    if t == "Vertex":
        return {"type": "Vertex", "rgb": [0.0, 0.0, 0.0]}
    return {"type": "Rgb", "rgb": _col(d)}


def _basic_material(d):
    t = d["type"]
    if t == "Lambertian":
        return {"type": "Lambertian", "albedo": _albedo(d["albedo"])}
    if t == "Gloss":
        return {"type": "Gloss", "albedo": _albedo(d["albedo"]), "reflectance": float(d["reflectance"]),
                "metalness": float(d.get("metalness", 0.0))}  # superset: stale scenes have no metalness
    if t == "Mirror":
        return {"type": "Mirror"}
    if t == "CookTorrance":
        return {"type": "CookTorrance", "albedo": {"type": "Rgb", "rgb": _col(d["albedo"])}, "roughness": float(d["roughness"])}
    raise ValueError("unknown material type %r" % t)


def _material(d):
    t = d["type"]
    if t == "Auto":
        return {"type": "Auto"}
    if t == "Fresnel":
        return {"type": "Fresnel", "refractive_index": float(d["refractive_index"]),
                "diffuse": _basic_material(d["diffuse"]), "specular": _basic_material(d["specular"])}
    return _basic_material(d)


def resolve_asset(path, asset_root):
    if os.path.isabs(path) or os.path.exists(path):
        return path
    if asset_root is not None:
        cand = os.path.join(asset_root, path)
        if os.path.exists(cand):
            return cand
    return path


def load_scene_yaml(yaml_path, asset_root=None, width=None, height=None):
    """SceneDescription::{camera,scene} (src/serde.rs:65-154).  Model paths are relative to the working directory
    in the reference; `asset_root` (default: the directory above the YAML's directory) is tried as well."""
    yaml_path = str(yaml_path)
    if asset_root is None:
        asset_root = str(Path(yaml_path).resolve().parent.parent)
    with open(yaml_path, "r") as fh:
        doc = yaml.safe_load(fh)
    cam = doc["camera"]
    camera = {
        "image_width": int(width if width is not None else cam["image_width"]),
        "image_height": int(height if height is not None else cam["image_height"]),
        "location": _vec(cam["location"]),
        "pitch": float(cam["orientation"]["pitch"]), "yaw": float(cam["orientation"]["yaw"]), "roll": float(cam["orientation"]["roll"]),
        "sensor_width": float(cam["sensor_width"]), "sensor_height": float(cam["sensor_height"]),
        "focal_length": float(cam["focal_length"]), "focus_distance": float(cam["focus_distance"]), "aperture": float(cam["aperture"]),
    }
    declared = {k: v["file"] for k, v in (doc.get("models") or {}).items()}
    models, loaded, objects = [], {}, []
    for o in doc.get("objects") or []:
        shp = o["shape"]
        if shp["type"] == "Sphere":
            objects.append({"shape": "Sphere", "center": _vec(shp["center"]), "radius": float(shp["radius"]),
                            "material": _material(o["material"])})
        elif shp["type"] == "Mesh":
            name = shp["model"]
            if name not in declared:
                raise ValueError("Attempt to load model %r before declaration" % name)
            if name not in loaded:  # src/model.rs:49-94
                fp = resolve_asset(declared[name], asset_root)
                ext = os.path.splitext(fp)[1].lower()
                if ext == ".obj":
                    ms = load_obj(fp)
                elif ext == ".ply":
                    ms = [load_ply(fp)]
                else:
                    raise ValueError("Unknown file extension: %s" % ext)
                loaded[name] = list(range(len(models), len(models) + len(ms)))
                models.extend(ms)
            rot = shp["rotation"]
            for ix in loaded[name]:  # one Object per sub-model, src/serde.rs:112-141
                mat = _material(o["material"])
                if mat["type"] == "Auto":
                    mat = models[ix]["material"] or {"type": "Lambertian", "albedo": {"type": "Rgb", "rgb": [1.0, 1.0, 1.0]}}
                objects.append({
                    "shape": "Mesh", "model": ix, "smooth_normals": bool(shp.get("smooth_normals", True)),
                    "translation": _vec(shp["translation"]),
                    "rotation": (float(rot["pitch"]), float(rot["yaw"]), float(rot["roll"])),
                    "scale": float(shp["scale"]), "material": mat,
                })
        else:
            raise ValueError("unknown shape type %r" % shp["type"])
    lights = []
    for l in doc.get("lights") or []:  # superset: `lights` may be absent (stale shipped scenes)
        g = l["geometry"]
        if g["type"] == "Point":
            lights.append({"geometry": "Point", "v": _vec(g), "radius": 0.0, "colour": _col(l["colour"]), "intensity": float(l["intensity"])})
        elif g["type"] == "Sphere":
            lights.append({"geometry": "Sphere", "v": _vec(g["center"]), "radius": float(g["radius"]),
                           "colour": _col(l["colour"]), "intensity": float(l["intensity"])})
        else:
            raise ValueError("unknown light geometry %r" % g["type"])
    sk = doc["skybox"]
    if sk["type"] == "Flat":
        sky = {"type": "Flat", "c0": _col(sk["colour"]), "c1": [0.0, 0.0, 0.0]}
    elif sk["type"] == "Gradient":
        sky = {"type": "Gradient", "c0": _col(sk["overhead_colour"]), "c1": _col(sk["horizon_colour"])}
    elif sk["type"] == "Hdri":
        w, h, tex = load_hdr(resolve_asset(sk["filename"], asset_root))
        sky = {"type": "Hdri", "c0": [0.0] * 3, "c1": [0.0] * 3, "width": w, "height": h, "texels": tex}
    else:
        raise ValueError("unknown skybox type %r" % sk["type"])
    return {"camera": camera, "models": models, "objects": objects, "lights": lights, "sky": sky}


def stress_scene(num_spheres=500, seed=0):
    """src/stress.rs:7-66 with a seeded generator (the reference's thread_rng is unseedable): SplitMix64 consumed in the reference's
    draw order -- centre x, y, z, radius, material choice, colour r, g, b, then the Gloss reflectance."""
    mask = (1 << 64) - 1
    state = [seed & mask]

    def unit():
        state[0] = (state[0] + 0x9E3779B97F4A7C15) & mask
        z = state[0]
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & mask
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & mask
        z ^= z >> 31
        return (z >> 11) * (1.0 / 9007199254740992.0)

    objects = []
    for _ in range(num_spheres):
        center = [unit() * 100.0 - 50.0, unit() * 100.0 - 50.0, unit() * 100.0]
        radius = unit() * 5.0
        choice = int(unit() * 3.0)
        colour = [unit(), unit(), unit()]
        if choice == 0:
            mat = {"type": "Gloss", "albedo": {"type": "Rgb", "rgb": colour}, "reflectance": 1.0 + unit() * 2.0, "metalness": 0.0}
        elif choice == 1:
            mat = {"type": "Lambertian", "albedo": {"type": "Rgb", "rgb": colour}}
        else:
            mat = {"type": "Mirror"}
        objects.append({"shape": "Sphere", "center": center, "radius": radius, "material": mat})
    camera = {"image_width": 720, "image_height": 480, "location": [0.0, -5.0, -13.0], "pitch": 0.0, "yaw": 0.0, "roll": -0.3,
              "sensor_width": 0.036, "sensor_height": 0.024, "focal_length": 0.05, "focus_distance": 10.0, "aperture": 8.0}
    sky = {"type": "Flat", "c0": [0.8, 0.8, 0.8], "c1": [0.0, 0.0, 0.0]}
    return {"camera": camera, "models": [], "objects": objects, "lights": [], "sky": sky}
