"""paths_b200 -- host-side mirror of the reference's render interface over the C ABI of libpaths_b200.so.

`Scene` stands for `SceneDescription::scene()` + `Scene::new` (src/serde.rs:81-154, src/scene.rs:143-170) and `Renderer` for
`Renderer` (src/renderer.rs:33-141): same method names and argument meaning, errors raised as `PathsError` where the
reference panics.  All compute happens in hand-written sm_100a kernels; there is no CPU path behind this module.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi
from ._capi import (TRACE_CAMERA, TRACE_FOUR_WIDE, TRACE_PLAIN, TRACE_TWO_WIDE, EXT_ENV_NEE, EXT_GGX, EXT_MICROFACET, EXT_MICROFACET_COSINE, EXT_NEE_ALL_LIGHTS, MISS, PRECISION_AUTO, PRECISION_F32, PRECISION_F64, PRECISION_MIXED, SHARD_SAMPLES, SHARD_TILES, Camera, PathsError,  # noqa: F401
                    RenderOpts, SceneInfo, Stats, build, check, lib)

__all__ = ["Scene", "Renderer", "RendererGroup", "comm_unique_id", "nccl_info", "Image", "Camera", "PathsError", "build", "lib", "camera_rays", "device_count", "MISS"]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def device_count():
    n = C.c_int32(0)
    st = lib().paths_device_count(C.byref(n))
    return n.value if st == _capi.OK else 0


def camera_from_dict(d):
    c = Camera()
    c.image_width, c.image_height = int(d["image_width"]), int(d["image_height"])
    c.location[:] = d["location"]
    c.pitch, c.yaw, c.roll = d["pitch"], d["yaw"], d["roll"]
    for k in ("sensor_width", "sensor_height", "focal_length", "focus_distance", "aperture"):
        setattr(c, k, d[k])
    return c


def _basic(d, out):
    out.type = _capi.MAT_TYPES[d["type"]]
    alb = d.get("albedo", {"type": "Rgb", "rgb": [0.0, 0.0, 0.0]})
    out.albedo_mode = 1 if alb["type"] == "Vertex" else 0
    out.albedo[:] = alb["rgb"]
    out.reflectance = d.get("reflectance", 0.0)
    out.metalness = d.get("metalness", 0.0)
    out.roughness = d.get("roughness", 0.0)


def scene_desc_from_dict(desc):
    """dict {models, objects, lights, sky} (flattened SceneDescription) -> (SceneDesc, keepalive)."""
    keep = []
    models = (_capi.ModelDesc * max(1, len(desc["models"])))()
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
    objects = (_capi.ObjectDesc * max(1, len(desc["objects"])))()
    for i, o in enumerate(desc["objects"]):
        od = objects[i]
        if o["shape"] == "Sphere":
            od.shape = 0
            od.center[:] = o["center"]
            od.radius = o["radius"]
            od.scale = 1.0
        else:
            od.shape = 1
            od.model = o["model"]
            od.smooth_normals = 1 if o.get("smooth_normals", True) else 0
            od.translation[:] = o["translation"]
            od.rot_pitch, od.rot_yaw, od.rot_roll = o["rotation"]
            od.scale = o["scale"]
        md = o["material"]
        if md["type"] == "Fresnel":
            od.material.m.type = _capi.MAT_TYPES["Fresnel"]
            od.material.refractive_index = md["refractive_index"]
            _basic(md["diffuse"], od.material.diffuse)
            _basic(md["specular"], od.material.specular)
        else:
            _basic(md, od.material.m)
            od.material.refractive_index = 1.0
    lights = (_capi.LightDesc * max(1, len(desc["lights"])))()
    for i, l in enumerate(desc["lights"]):
        lights[i].geometry = 0 if l["geometry"] == "Point" else 1
        lights[i].v[:] = l["v"]
        lights[i].radius = l.get("radius", 0.0)
        lights[i].colour[:] = l["colour"]
        lights[i].intensity = l["intensity"]
    sd = _capi.SceneDesc()
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


class Image:
    """Image {width, height, pixels} (src/camera.rs:6-10): row-major linear-radiance means, index x + y * width."""

    def __init__(self, width, height, pixels):
        self.width, self.height, self.pixels = width, height, pixels  # pixels: float32 [H, W, 3]


class Scene:
    """A scene staged for rendering.  Build from a YAML file (`Scene.load`) or a flattened description (`Scene(desc)`)."""

    def __init__(self, desc=None, _handle=None, camera=None):
        self._h = None
        self.camera = camera
        if _handle is not None:
            self._h = _handle
            return
        sd, keep = scene_desc_from_dict(desc)
        h = C.c_void_p()
        check(lib().paths_scene_create(C.byref(sd), C.byref(h)))
        self._h = h
        if camera is None and desc.get("camera"):
            self.camera = camera_from_dict(desc["camera"])

    @classmethod
    def load(cls, yaml_path, asset_root=None, width=0, height=0):
        """serde_yaml::from_reader + SceneDescription::{camera, scene} (src/main.rs:43-53)."""
        h = C.c_void_p()
        cam = Camera()
        check(lib().paths_scene_load_yaml(str(yaml_path).encode(), asset_root.encode() if asset_root else None, int(width), int(height),
                                          C.byref(h), C.byref(cam)))
        return cls(_handle=h, camera=cam)

    @classmethod
    def stress(cls, num_spheres=500, seed=0):
        """generate_stress_scene (src/stress.rs:7-66), seeded; what the reference renders when started without a scene file."""
        h = C.c_void_p()
        cam = Camera()
        check(lib().paths_scene_create_stress(int(num_spheres), int(seed), C.byref(h), C.byref(cam)))
        return cls(_handle=h, camera=cam)

    def close(self):
        if self._h:
            lib().paths_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def info(self):
        si = SceneInfo()
        check(lib().paths_scene_get_info(self._h, C.byref(si)))
        return si.as_dict()

    def build_bvh(self, reference=False, device=False):
        """construct_bvh_aac (src/bvh.rs:160-206) when reference=True, else the production tree: the host SAH builder, or with
        device=True the GPU morton builder (runs at staging time on each device); returns `info`."""
        check(lib().paths_scene_build_bvh(self._h, 2 if device else (1 if reference else 0)))
        return self.info

    def check_bvh(self, device=0):
        """Walks the production tree staged on `device` on the host; returns the number of structural violations."""
        bad = C.c_uint32(0)
        check(lib().paths_scene_check_bvh(self._h, device, C.byref(bad)))
        return bad.value

    def prim_entity(self, prim):
        il, en, fa = C.c_int32(), C.c_uint32(), C.c_uint32()
        check(lib().paths_scene_prim_entity(self._h, int(prim), C.byref(il), C.byref(en), C.byref(fa)))
        return bool(il.value), en.value, fa.value

    def prim_geometry(self, prim):
        """(is_triangle, 12 doubles): a, b, c, n of a triangle or centre, radius of a sphere, after instancing."""
        out = np.zeros(12)
        tri = C.c_int32()
        check(lib().paths_scene_prim_geometry(self._h, int(prim), _p(out), C.byref(tri)))
        return bool(tri.value), out

    def find_intersection(self, origins, dirs, precision=PRECISION_F64, device=0, want_bary=False):
        """BVH::find_intersection (src/bvh.rs:78-140) over a batch of rays -> (prim_id, t[, bary])."""
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        prim = np.empty(n, dtype=np.uint32)
        t = np.empty(n, dtype=np.float64)
        bary = np.empty((n, 3), dtype=np.float64) if want_bary else None
        check(lib().paths_trace_closest(self._h, device, _p(o), _p(d), n, _p(prim), _p(t), _p(bary) if want_bary else None, precision))
        return (prim, t, bary) if want_bary else (prim, t)

    STATIC_MAX_RAYS = 65536  # PATHS_STATIC_MAX_RAYS

    def find_intersection_static(self, origins, dirs, device=0, want_bary=False):
        """paths_trace_closest_static: find_intersection(precision=32) with every byte of device scratch in a static device array of the
        library (no allocation; at most STATIC_MAX_RAYS rays; the scene must already be staged on the device).  No reference counterpart."""
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        prim = np.empty(n, dtype=np.uint32)
        t = np.empty(n, dtype=np.float64)
        bary = np.empty((n, 3), dtype=np.float64) if want_bary else None
        check(lib().paths_trace_closest_static(self._h, device, _p(o), _p(d), n, _p(prim), _p(t), _p(bary) if want_bary else None))
        return (prim, t, bary) if want_bary else (prim, t)

    def occluded(self, origins, dirs, light=None, precision=PRECISION_F32, device=0):
        """The shadow query of trace_ray (src/trace.rs:61-66) over a batch of rays, through the visibility kernel a render launches.
        light: index (or per-ray indices) of the light the rays were sampled toward, None = every hit occludes.  -> uint8 flags."""
        o = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        lt = None
        if light is not None:
            lt = np.ascontiguousarray(np.broadcast_to(np.asarray(light, dtype=np.uint32), (n,)))
        occ = np.empty(n, dtype=np.uint8)
        check(lib().paths_trace_occluded(self._h, device, _p(o), _p(d), _p(lt) if lt is not None else None, n, _p(occ), precision))
        return occ


class Renderer:
    """Renderer (src/renderer.rs:33-141) on one GPU.

    `num_workers` of the reference has no meaning here; `rank` / `world` / `shard_mode` select this process's share of the
    frame (samples or 32x32 tiles) when several GPUs render one image.
    """

    def __init__(self, camera, scene, device=0, precision=PRECISION_AUTO, seed=1, rank=0, world=1, shard_mode=SHARD_SAMPLES,
                 max_bounces=10, samples_per_pass=0, time_kernels=False, extensions=0):
        self._h = None
        self.scene = scene
        self.camera = camera if isinstance(camera, Camera) else camera_from_dict(camera)
        o = RenderOpts()
        lib().paths_render_opts_default(C.byref(o))
        o.device, o.precision, o.seed, o.rank, o.world, o.shard_mode = device, precision, seed, rank, world, shard_mode
        o.max_bounces, o.samples_per_pass, o.time_kernels = max_bounces, samples_per_pass, 1 if time_kernels else 0
        o.extensions = extensions
        h = C.c_void_p()
        check(lib().paths_renderer_create(scene._h, C.byref(self.camera), C.byref(o), C.byref(h)))
        self._h = h
        self.width, self.height = self.camera.image_width, self.camera.image_height

    # -- renderer.rs:79-110: the reference pumps request/result queues once per UI tick; here one call advances every pixel
    def render_samples(self, spp, wait=True):
        check((lib().paths_renderer_render if wait else lib().paths_renderer_render_async)(self._h, int(spp)))

    def fill_request_queue(self, spp=1):
        self.render_samples(spp, wait=False)

    def drain_result_queue(self):
        check(lib().paths_renderer_sync(self._h))

    def render(self):
        """renderer.rs:71-73 -> Image of per-pixel means."""
        px = np.empty((self.height, self.width, 3), dtype=np.float32)
        check(lib().paths_renderer_image_f32(self._h, _p(px)))
        return Image(self.width, self.height, px)

    def render_rgb24(self):
        """Colour::to_bytes packed as RGB24, pitch 3W (colour.rs:30-36,62-74; main.rs:114-120)."""
        px = np.empty((self.height, self.width, 3), dtype=np.uint8)
        check(lib().paths_renderer_image_rgb24(self._h, _p(px)))
        return px

    def num_rays_cast(self):
        """renderer.rs:75-77: counts camera samples."""
        return self.stats()["samples"]

    def stats(self):
        st = Stats()
        check(lib().paths_renderer_get_stats(self._h, C.byref(st)))
        return st.as_dict()

    def sums(self):
        """Estimator state (pixels.rs:6-31): float64 [H, W, 4] = sum r, g, b and sample count."""
        s = np.empty((self.height, self.width, 4), dtype=np.float64)
        check(lib().paths_renderer_get_sums(self._h, _p(s)))
        return s

    def add_sums(self, sums4):
        s = np.ascontiguousarray(sums4, dtype=np.float64)
        assert s.size == self.width * self.height * 4
        check(lib().paths_renderer_add_sums(self._h, _p(s)))

    def device_sums(self):
        """(device pointer, element count, element bytes) of the estimator buffer, for in-place NCCL collectives."""
        ptr, n, eb = C.c_void_p(), C.c_size_t(), C.c_int32()
        check(lib().paths_renderer_device_sums(self._h, C.byref(ptr), C.byref(n), C.byref(eb)))
        return ptr.value, n.value, eb.value

    def cuda_stream(self):
        """cudaStream_t (as an int) all kernels of this renderer run on."""
        s = C.c_void_p()
        check(lib().paths_renderer_cuda_stream(self._h, C.byref(s)))
        return s.value or 0

    def set_camera(self, location, orientation):
        """renderer.rs:112-120: orientation is a row-major 3x3 rotation; starts a new epoch."""
        loc = np.ascontiguousarray(location, dtype=np.float64)
        rot = np.ascontiguousarray(orientation, dtype=np.float64).reshape(9)
        check(lib().paths_renderer_set_camera(self._h, _p(loc), _p(rot)))

    def reset(self):
        check(lib().paths_renderer_reset(self._h))

    # -- the interactive front half (renderer.rs:143-192)
    def request_preview(self):
        """renderer.rs:152-164: one sample for every 6th pixel; `render()` fills unsampled pixels from that grid (pixels.rs:53-79)."""
        check(lib().paths_renderer_render_preview(self._h))

    def render_columns(self, n_requests):
        """The next `n_requests` column requests of next_request (renderer.rs:166-192), centre outwards."""
        check(lib().paths_renderer_render_columns(self._h, int(n_requests)))

    def progress(self):
        """(samples per pixel completed, columns done in the sweep in progress)."""
        a, b = C.c_uint32(), C.c_uint32()
        check(lib().paths_renderer_get_progress(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def save_checkpoint(self, path):
        check(lib().paths_renderer_save_checkpoint(self._h, str(path).encode()))

    def load_checkpoint(self, path):
        check(lib().paths_renderer_load_checkpoint(self._h, str(path).encode()))

    # -- several GPUs, one frame: the combine of partial results lives inside Renderer in the reference too (renderer.rs:93-105)
    def attach_comm(self, unique_id):
        """Joins this renderer's rank to the frame's NCCL communicator (collective over all `world` ranks).  unique_id: the 128 bytes
        rank 0 got from comm_unique_id(), handed to every rank out of band."""
        _preload_nccl()
        buf = (C.c_uint8 * _capi.COMM_ID_BYTES).from_buffer_copy(bytes(unique_id))
        check(lib().paths_renderer_attach_comm(self._h, buf))

    def combine(self, root=0, wait=True):
        """Combines the ranks' partial estimators on `root` over NCCL (sample shards: reduce; tile shards: send / receive of owned tiles).
        On the root, render() / render_rgb24() / sums() then show the combined frame until new samples are rendered."""
        check((lib().paths_renderer_combine if wait else lib().paths_renderer_combine_async)(self._h, int(root)))

    def shutdown(self):
        if self._h:
            lib().paths_renderer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.shutdown()
        except Exception:
            pass


_nccl_preloaded = False


def _preload_nccl():
    """The library binds whatever libnccl.so.2 the process already has.  In a Python process that may import PyTorch LATER, the copy
    bound first must be the one PyTorch ships (an older system NCCL lacks symbols libtorch_cuda needs): map it before the first NCCL call."""
    global _nccl_preloaded
    if _nccl_preloaded:
        return
    _nccl_preloaded = True
    import importlib.util
    import os

    try:
        spec = importlib.util.find_spec("nvidia.nccl")
    except (ImportError, ValueError):
        spec = None
    if spec and spec.submodule_search_locations:
        for base in spec.submodule_search_locations:
            path = os.path.join(base, "lib", "libnccl.so.2")
            if os.path.exists(path):
                try:
                    C.CDLL(path, mode=C.RTLD_GLOBAL)
                except OSError:
                    pass
                return


def nccl_info():
    """Which NCCL the library bound at run time (version and file), or why none could be."""
    _preload_nccl()
    return lib().paths_nccl_info().decode("utf-8", "replace")


def comm_unique_id():
    """ncclGetUniqueId as 128 bytes (rank 0 calls it and broadcasts the result)."""
    _preload_nccl()
    buf = (C.c_uint8 * _capi.COMM_ID_BYTES)()
    check(lib().paths_comm_unique_id(buf))
    return bytes(buf)


class RendererGroup:
    """One frame rendered by several GPUs of this process: one renderer, stream and NCCL rank per device, owned by the library.
    render_samples(spp) advances the frame by spp samples per pixel in total and combines on the first device."""

    def __init__(self, camera, scene, devices, precision=PRECISION_AUTO, seed=1, shard_mode=SHARD_SAMPLES, max_bounces=10,
                 samples_per_pass=0, extensions=0):
        self._h = None
        self.scene = scene
        self.camera = camera if isinstance(camera, Camera) else camera_from_dict(camera)
        devices = list(range(devices)) if isinstance(devices, int) else list(devices)
        o = RenderOpts()
        lib().paths_render_opts_default(C.byref(o))
        o.precision, o.seed, o.shard_mode, o.max_bounces, o.samples_per_pass, o.extensions = precision, seed, shard_mode, max_bounces, samples_per_pass, extensions
        ids = (C.c_int32 * len(devices))(*devices)
        if len(devices) > 1:
            _preload_nccl()
        h = C.c_void_p()
        check(lib().paths_group_create(scene._h, C.byref(self.camera), C.byref(o), len(devices), ids, C.byref(h)))
        self._h = h
        self.n_devices = len(devices)
        self.width, self.height = self.camera.image_width, self.camera.image_height

    def render_samples(self, spp):
        check(lib().paths_group_render(self._h, int(spp)))

    def render(self):
        px = np.empty((self.height, self.width, 3), dtype=np.float32)
        check(lib().paths_group_image_f32(self._h, _p(px)))
        return Image(self.width, self.height, px)

    def render_rgb24(self):
        px = np.empty((self.height, self.width, 3), dtype=np.uint8)
        check(lib().paths_group_image_rgb24(self._h, _p(px)))
        return px

    def sums(self):
        s = np.empty((self.height, self.width, 4), dtype=np.float64)
        check(lib().paths_group_get_sums(self._h, _p(s)))
        return s

    def stats(self, rank=None):
        st = Stats()
        if rank is None:
            check(lib().paths_group_get_stats(self._h, C.byref(st)))
        else:
            r = C.c_void_p()
            check(lib().paths_group_renderer(self._h, int(rank), C.byref(r)))
            check(lib().paths_renderer_get_stats(r, C.byref(st)))
        return st.as_dict()

    def set_camera(self, location, orientation):
        loc = np.ascontiguousarray(location, dtype=np.float64)
        rot = np.ascontiguousarray(orientation, dtype=np.float64).reshape(9)
        check(lib().paths_group_set_camera(self._h, _p(loc), _p(rot)))

    def reset(self):
        check(lib().paths_group_reset(self._h))

    def shutdown(self):
        if self._h:
            lib().paths_group_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.shutdown()
        except Exception:
            pass


def camera_rays(camera, seed, sample, x0=0, y0=0, w=None, h=None, precision=PRECISION_F64, device=0):
    """Camera::get_ray_for_pixel for sample `sample` of every pixel of a rectangle (camera.rs:47-94, worker.rs:58-87)."""
    cam = camera if isinstance(camera, Camera) else camera_from_dict(camera)
    w = cam.image_width if w is None else w
    h = cam.image_height if h is None else h
    n = w * h
    o, d, wt = np.empty((n, 3)), np.empty((n, 3)), np.empty(n)
    check(lib().paths_camera_rays(C.byref(cam), device, seed, sample, x0, y0, w, h, _p(o), _p(d), _p(wt), precision))
    return o, d, wt
