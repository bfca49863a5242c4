"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE ONLY).

The oracle (oracle/oracle.cpp) is a CPU f64 restatement of the reference renderer
(rynorris/paths).  It is the checker for tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  It is never imported by the
product package (paths_b200).  Parity status: pinned only by the reference's own
unit-test vectors (see oracle/oracle.h); everything else is "parity unpinned".
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

from . import scene_io  # noqa: F401

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "liboracle.so"

MAT_TYPES = {"Lambertian": 0, "Mirror": 1, "Gloss": 2, "CookTorrance": 3, "Fresnel": 4}
MISS = 0xFFFFFFFF


def build(force=False):
    src = [_HERE / "oracle.cpp", _HERE / "oracle.h"]
    if force or not _LIB_PATH.exists() or any(s.stat().st_mtime > _LIB_PATH.stat().st_mtime for s in src):
        subprocess.check_call(["make", "-C", str(_HERE), "-s"] + (["-B"] if force else []))
    return _LIB_PATH


class ModelDesc(C.Structure):
    _fields_ = [("n_vertices", C.c_uint32), ("n_faces", C.c_uint32), ("vertices", C.POINTER(C.c_double)),
                ("faces", C.POINTER(C.c_uint32)), ("vertex_colours", C.POINTER(C.c_double))]


class BasicMaterial(C.Structure):
    _fields_ = [("type", C.c_int32), ("albedo_mode", C.c_int32), ("albedo", C.c_double * 3), ("reflectance", C.c_double),
                ("metalness", C.c_double), ("roughness", C.c_double)]


class MaterialDesc(C.Structure):
    _fields_ = [("m", BasicMaterial), ("refractive_index", C.c_double), ("diffuse", BasicMaterial), ("specular", BasicMaterial)]


class ObjectDesc(C.Structure):
    _fields_ = [("shape", C.c_int32), ("model", C.c_int32), ("smooth_normals", C.c_int32), ("_pad", C.c_int32),
                ("center", C.c_double * 3), ("radius", C.c_double), ("translation", C.c_double * 3),
                ("rot_pitch", C.c_double), ("rot_yaw", C.c_double), ("rot_roll", C.c_double), ("scale", C.c_double),
                ("material", MaterialDesc)]


class LightDesc(C.Structure):
    _fields_ = [("geometry", C.c_int32), ("_pad", C.c_int32), ("v", C.c_double * 3), ("radius", C.c_double),
                ("colour", C.c_double * 3), ("intensity", C.c_double)]


class SkyDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("width", C.c_uint32), ("height", C.c_uint32), ("_pad", C.c_uint32),
                ("c0", C.c_double * 3), ("c1", C.c_double * 3), ("texels", C.POINTER(C.c_float))]


class SceneDesc(C.Structure):
    _fields_ = [("n_models", C.c_uint32), ("n_objects", C.c_uint32), ("n_lights", C.c_uint32), ("_pad", C.c_uint32),
                ("models", C.POINTER(ModelDesc)), ("objects", C.POINTER(ObjectDesc)), ("lights", C.POINTER(LightDesc)),
                ("sky", SkyDesc)]


class CameraDesc(C.Structure):
    _fields_ = [("image_width", C.c_uint32), ("image_height", C.c_uint32), ("location", C.c_double * 3),
                ("pitch", C.c_double), ("yaw", C.c_double), ("roll", C.c_double), ("sensor_width", C.c_double),
                ("sensor_height", C.c_double), ("focal_length", C.c_double), ("focus_distance", C.c_double), ("aperture", C.c_double)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("samples", "extend_rays", "shadow_rays", "box_tests", "tri_tests", "sphere_tests",
                                          "segments", "bad_samples")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class PrimInfo(C.Structure):
    _fields_ = [("type", C.c_int32), ("is_light", C.c_int32), ("entity", C.c_uint32), ("face", C.c_uint32), ("v", C.c_double * 9),
                ("normal", C.c_double * 3), ("aabb_min", C.c_double * 3), ("aabb_max", C.c_double * 3)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(_LIB_PATH))
        L.orc_scene_create.restype = C.c_void_p
        L.orc_scene_create.argtypes = [C.POINTER(SceneDesc)]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_last_error.restype = C.c_char_p
        for f in ("orc_scene_num_prims", "orc_scene_num_nodes", "orc_scene_bvh_depth"):
            getattr(L, f).restype = C.c_uint32
            getattr(L, f).argtypes = [C.c_void_p]
        L.orc_scene_prim_info.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(PrimInfo)]
        L.orc_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(Counters)]
        L.orc_intersect_prim.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
        L.orc_camera_rays.argtypes = [C.POINTER(CameraDesc), C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                      C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_render.argtypes = [C.c_void_p, C.POINTER(CameraDesc), C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32,
                                 C.c_int, C.c_void_p, C.c_void_p, C.POINTER(Counters)]
        L.orc_trace_pixel.argtypes = [C.c_void_p, C.POINTER(CameraDesc), C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_permute.restype = C.c_uint32
        L.orc_permute.argtypes = [C.c_uint32] * 3
        L.orc_rand_float.restype = C.c_double
        L.orc_rand_float.argtypes = [C.c_uint32] * 2
        L.orc_cmj.argtypes = [C.c_uint32] * 4 + [C.c_void_p]
        L.orc_cmj_disk.argtypes = [C.c_uint32] * 4 + [C.c_void_p]
        L.orc_morton_code.restype = C.c_uint64
        L.orc_morton_code.argtypes = [C.c_uint16] * 4
        L.orc_matmul3.argtypes = [C.c_void_p] * 3
        L.orc_rotation.argtypes = [C.c_double] * 3 + [C.c_void_p]
        L.orc_switch_basis.argtypes = [C.c_void_p] * 3
        L.orc_to_bytes.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_philox4x32_10.argtypes = [C.c_void_p] * 3
        L.orc_pixel_grid.argtypes = [C.c_uint32] * 4 + [C.c_void_p]
        L.orc_sample_schedule.argtypes = [C.c_uint32, C.c_void_p]
        L.orc_sky.argtypes = [C.c_void_p] * 3
        L.orc_material_brdf.argtypes = [C.POINTER(MaterialDesc)] + [C.c_void_p] * 4
        L.orc_material_sample.restype = C.c_int
        L.orc_material_sample.argtypes = [C.POINTER(MaterialDesc)] + [C.c_void_p] * 4 + [C.POINTER(C.c_double), C.c_void_p]
        L.orc_light_sample.argtypes = [C.POINTER(LightDesc), C.c_void_p, C.c_double, C.c_double, C.c_void_p, C.POINTER(C.c_double)]
        L.orc_ccrf.restype = C.c_uint32
        L.orc_ccrf.argtypes = [C.c_uint32]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _basic(d, out):
    out.type = MAT_TYPES[d["type"]]
    alb = d.get("albedo", {"type": "Rgb", "rgb": [0.0, 0.0, 0.0]})
    out.albedo_mode = 1 if alb["type"] == "Vertex" else 0
    out.albedo[:] = alb["rgb"]
    out.reflectance = d.get("reflectance", 0.0)
    out.metalness = d.get("metalness", 0.0)
    out.roughness = d.get("roughness", 0.0)


def material_desc(d):
    m = MaterialDesc()
    if d["type"] == "Fresnel":
        m.m.type = MAT_TYPES["Fresnel"]
        m.refractive_index = d["refractive_index"]
        _basic(d["diffuse"], m.diffuse)
        _basic(d["specular"], m.specular)
    else:
        _basic(d, m.m)
        m.refractive_index = 1.0
    return m


def camera_desc(cam):
    c = CameraDesc()
    c.image_width, c.image_height = cam["image_width"], cam["image_height"]
    c.location[:] = cam["location"]
    c.pitch, c.yaw, c.roll = cam["pitch"], cam["yaw"], cam["roll"]
    for k in ("sensor_width", "sensor_height", "focal_length", "focus_distance", "aperture"):
        setattr(c, k, cam[k])
    return c


def light_desc(l):
    d = LightDesc()
    d.geometry = 0 if l["geometry"] == "Point" else 1
    d.v[:] = l["v"]
    d.radius = l.get("radius", 0.0)
    d.colour[:] = l["colour"]
    d.intensity = l["intensity"]
    return d


def build_scene_desc(desc):
    """dict (oracle.scene_io format) -> (SceneDesc, keepalive list)."""
    keep = []
    models = (ModelDesc * max(1, len(desc["models"])))()
    for i, m in enumerate(desc["models"]):
        v = np.ascontiguousarray(m["vertices"], dtype=np.float64)
        f = np.ascontiguousarray(m["faces"], dtype=np.uint32)
        keep += [v, f]
        models[i].n_vertices, models[i].n_faces = v.shape[0], f.shape[0]
        models[i].vertices = v.ctypes.data_as(C.POINTER(C.c_double))
        models[i].faces = f.ctypes.data_as(C.POINTER(C.c_uint32))
        if m.get("vertex_colours") is not None:
            vc = np.ascontiguousarray(m["vertex_colours"], dtype=np.float64)
            keep.append(vc)
            models[i].vertex_colours = vc.ctypes.data_as(C.POINTER(C.c_double))
    objects = (ObjectDesc * max(1, len(desc["objects"])))()
    for i, o in enumerate(desc["objects"]):
        od = objects[i]
        if o["shape"] == "Sphere":
            od.shape = 0
            od.center[:] = o["center"]
            od.radius = o["radius"]
            od.scale = 1.0
            od.smooth_normals = 0
        else:
            od.shape = 1
            od.model = o["model"]
            od.smooth_normals = 1 if o.get("smooth_normals", True) else 0
            od.translation[:] = o["translation"]
            od.rot_pitch, od.rot_yaw, od.rot_roll = o["rotation"]
            od.scale = o["scale"]
        od.material = material_desc(o["material"])
    lights = (LightDesc * max(1, len(desc["lights"])))()
    for i, l in enumerate(desc["lights"]):
        lights[i] = light_desc(l)
    sd = SceneDesc()
    sd.n_models, sd.n_objects, sd.n_lights = len(desc["models"]), len(desc["objects"]), len(desc["lights"])
    sd.models, sd.objects, sd.lights = models, objects, lights
    sky = desc["sky"]
    sd.sky.type = {"Flat": 0, "Gradient": 1, "Hdri": 2}[sky["type"]]
    sd.sky.c0[:] = sky["c0"]
    sd.sky.c1[:] = sky["c1"]
    if sky["type"] == "Hdri":
        tex = np.ascontiguousarray(sky["texels"], dtype=np.float32)
        keep.append(tex)
        sd.sky.width, sd.sky.height = sky["width"], sky["height"]
        sd.sky.texels = tex.ctypes.data_as(C.POINTER(C.c_float))
    keep += [models, objects, lights]
    return sd, keep


class Scene:
    """Scene::new (src/scene.rs:143-170) + AAC BVH (src/bvh.rs:160-206) on the CPU."""

    def __init__(self, desc):
        L = lib()
        sd, keep = build_scene_desc(desc)
        self._h = L.orc_scene_create(C.byref(sd))
        if not self._h:
            raise RuntimeError("oracle scene: " + L.orc_last_error().decode())
        self.desc = desc

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_scene_destroy(self._h)
            self._h = None

    @property
    def num_prims(self):
        return lib().orc_scene_num_prims(self._h)

    @property
    def num_nodes(self):
        return lib().orc_scene_num_nodes(self._h)

    @property
    def bvh_depth(self):
        return lib().orc_scene_bvh_depth(self._h)

    def prim_info(self, prim):
        pi = PrimInfo()
        if lib().orc_scene_prim_info(self._h, prim, C.byref(pi)) != 0:
            raise IndexError(prim)
        return pi

    def trace_closest(self, origins, dirs, want_bary=False, counters=None):
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        prim = np.empty(n, dtype=np.uint32)
        t = np.empty(n, dtype=np.float64)
        bary = np.empty((n, 3), dtype=np.float64) if want_bary else None
        ctr = counters if counters is not None else Counters()
        lib().orc_trace_closest(self._h, _p(o), _p(d), n, _p(prim), _p(t), _p(bary) if want_bary else None, C.byref(ctr))
        return (prim, t, bary) if want_bary else (prim, t)

    def intersect_prim(self, prim, o, d):
        o = np.ascontiguousarray(o, dtype=np.float64)
        d = np.ascontiguousarray(d, dtype=np.float64)
        t = C.c_double()
        hit = lib().orc_intersect_prim(self._h, int(prim), _p(o), _p(d), C.byref(t))
        return t.value if hit else None

    def render(self, camera, seed, s0, s1, shard_mode=0, rank=0, world=1, threads=0, sums=None, counts=None, counters=None):
        cd = camera_desc(camera)
        W, H = camera["image_width"], camera["image_height"]
        if sums is None:
            sums = np.zeros((H, W, 3), dtype=np.float64)
        if counts is None:
            counts = np.zeros((H, W), dtype=np.uint32)
        ctr = counters if counters is not None else Counters()
        lib().orc_render(self._h, C.byref(cd), seed, s0, s1, shard_mode, rank, world, threads, _p(sums), _p(counts), C.byref(ctr))
        return sums, counts, ctr

    def trace_pixel(self, camera, seed, x, y, sample):
        cd = camera_desc(camera)
        out = np.zeros(3, dtype=np.float64)
        lib().orc_trace_pixel(self._h, C.byref(cd), seed, x, y, sample, _p(out))
        return out

    def sky(self, direction):
        d = np.ascontiguousarray(direction, dtype=np.float64)
        out = np.zeros(3)
        lib().orc_sky(self._h, _p(d), _p(out))
        return out


def camera_rays(camera, seed, sample, x0=0, y0=0, w=None, h=None):
    cd = camera_desc(camera)
    w = camera["image_width"] if w is None else w
    h = camera["image_height"] if h is None else h
    n = w * h
    o = np.empty((n, 3))
    d = np.empty((n, 3))
    wt = np.empty(n)
    lib().orc_camera_rays(C.byref(cd), seed, sample, x0, y0, w, h, _p(o), _p(d), _p(wt))
    return o, d, wt


def to_bytes(rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.float64)
    out = np.empty(rgb.shape, dtype=np.uint8)
    lib().orc_to_bytes(_p(rgb), rgb.size // 3, _p(out))
    return out


def load_scene(yaml_path, **kw):
    desc = scene_io.load_scene_yaml(yaml_path, **kw)
    return Scene(desc), desc["camera"]
