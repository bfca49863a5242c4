"""ctypes declarations of include/paths_b200.h (the C ABI of libpaths_b200.so).

Loading never falls back: if the shared library is missing the import fails, and every compute entry point
returns PATHS_ERR_CUDA when no CUDA device is present.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import os

_PKG = Path(__file__).resolve().parent
# PATHS_B200_LIB points the mirror at another build of the same ABI (the debug-assert build, a variant under ab/); never at a fallback
LIB_PATH = Path(os.environ["PATHS_B200_LIB"]).resolve() if os.environ.get("PATHS_B200_LIB") else _PKG / "lib" / "libpaths_b200.so"

OK, ERR_INVALID_ARG, ERR_IO, ERR_PARSE, ERR_CUDA, ERR_UNSUPPORTED, ERR_INTERNAL, ERR_NCCL = range(8)
COMM_ID_BYTES = 128
MISS = 0xFFFFFFFF
PRECISION_AUTO, PRECISION_F32, PRECISION_MIXED, PRECISION_F64 = 0, 32, 48, 64
TRACE_CAMERA, TRACE_TWO_WIDE, TRACE_PLAIN, TRACE_FOUR_WIDE = 33, 34, 35, 36  # f32 kernels the per-ray seam can be pointed at
SHARD_SAMPLES, SHARD_TILES = 0, 1
EXT_ENV_NEE, EXT_MICROFACET, EXT_MICROFACET_COSINE, EXT_GGX, EXT_NEE_ALL_LIGHTS = 1, 2, 4, 8, 16
MAT_TYPES = {"Lambertian": 0, "Mirror": 1, "Gloss": 2, "CookTorrance": 3, "Fresnel": 4}


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


class Camera(C.Structure):
    _fields_ = [("image_width", C.c_uint32), ("image_height", C.c_uint32), ("location", C.c_double * 3),
                ("pitch", C.c_double), ("yaw", C.c_double), ("roll", C.c_double), ("sensor_width", C.c_double),
                ("sensor_height", C.c_double), ("focal_length", C.c_double), ("focus_distance", C.c_double), ("aperture", C.c_double)]

    def as_dict(self):
        return {"image_width": self.image_width, "image_height": self.image_height, "location": list(self.location),
                "pitch": self.pitch, "yaw": self.yaw, "roll": self.roll, "sensor_width": self.sensor_width,
                "sensor_height": self.sensor_height, "focal_length": self.focal_length, "focus_distance": self.focus_distance,
                "aperture": self.aperture}


class RenderOpts(C.Structure):
    _fields_ = [("device", C.c_int32), ("precision", C.c_int32), ("shard_mode", C.c_int32), ("rank", C.c_uint32), ("world", C.c_uint32),
                ("seed", C.c_uint64), ("max_bounces", C.c_uint32), ("samples_per_pass", C.c_uint32), ("time_kernels", C.c_int32),
                ("extensions", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("extend_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("bad_samples", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("passes", C.c_uint64), ("ms_raygen", C.c_double), ("ms_extend", C.c_double),
                ("ms_shade", C.c_double), ("ms_shadow", C.c_double), ("ms_accumulate", C.c_double), ("ms_total", C.c_double),
                ("n_extend_launches", C.c_uint64), ("n_shadow_launches", C.c_uint64),
                ("ms_device", C.c_double), ("stack_overflows", C.c_uint64), ("assert_failures", C.c_uint64), ("extend_kernel", C.c_int32),
                ("camera_kernel", C.c_int32), ("shade_kernel", C.c_int32), ("pass_contexts", C.c_int32), ("debug_asserts", C.c_int32), ("_pad", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class SceneInfo(C.Structure):
    _fields_ = [("n_prims", C.c_uint32), ("n_triangles", C.c_uint32), ("n_spheres", C.c_uint32), ("n_objects", C.c_uint32),
                ("n_lights", C.c_uint32), ("ref_nodes", C.c_uint32), ("ref_depth", C.c_uint32), ("fast_nodes", C.c_uint32),
                ("fast_depth", C.c_uint32), ("hidden_prims", C.c_uint32), ("auto_precision", C.c_int32), ("bounds_min", C.c_double * 3),
                ("bounds_max", C.c_double * 3), ("build_ms_ref", C.c_double), ("build_ms_fast", C.c_double), ("build_ms_device", C.c_double),
                ("bvh_builder", C.c_uint32), ("device_nodes2", C.c_uint32), ("device_nodes4", C.c_uint32), ("device_depth2", C.c_uint32),
                ("device_depth4", C.c_uint32), ("giant_prims", C.c_uint32)]

    def as_dict(self):
        out = {}
        for n, _ in self._fields_:
            v = getattr(self, n)
            out[n] = list(v) if hasattr(v, "__len__") else v
        return out


# every symbol include/paths_b200.h declares: name -> (restype, argtypes)
VP = C.c_void_p
SIGNATURES = {
    "paths_last_error": (C.c_char_p, []),
    "paths_version": (C.c_char_p, []),
    "paths_device_count": (C.c_int, [C.POINTER(C.c_int32)]),
    "paths_scene_create": (C.c_int, [C.POINTER(SceneDesc), C.POINTER(VP)]),
    "paths_scene_load_yaml": (C.c_int, [C.c_char_p, C.c_char_p, C.c_uint32, C.c_uint32, C.POINTER(VP), C.POINTER(Camera)]),
    "paths_scene_create_stress": (C.c_int, [C.c_uint32, C.c_uint64, C.POINTER(VP), C.POINTER(Camera)]),
    "paths_scene_destroy": (None, [VP]),
    "paths_scene_build_bvh": (C.c_int, [VP, C.c_int32]),
    "paths_scene_get_info": (C.c_int, [VP, C.POINTER(SceneInfo)]),
    "paths_scene_check_bvh": (C.c_int, [VP, C.c_int32, C.POINTER(C.c_uint32)]),
    "paths_scene_prim_geometry": (C.c_int, [VP, C.c_uint32, VP, C.POINTER(C.c_int32)]),
    "paths_scene_prim_entity": (C.c_int, [VP, C.c_uint32, C.POINTER(C.c_int32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "paths_trace_closest": (C.c_int, [VP, C.c_int32, VP, VP, C.c_size_t, VP, VP, VP, C.c_int32]),
    "paths_trace_closest_static": (C.c_int, [VP, C.c_int32, VP, VP, C.c_size_t, VP, VP, VP]),
    "paths_trace_occluded": (C.c_int, [VP, C.c_int32, VP, VP, VP, C.c_size_t, VP, C.c_int32]),
    "paths_render_opts_default": (None, [C.POINTER(RenderOpts)]),
    "paths_renderer_create": (C.c_int, [VP, C.POINTER(Camera), C.POINTER(RenderOpts), C.POINTER(VP)]),
    "paths_renderer_render": (C.c_int, [VP, C.c_uint32]),
    "paths_renderer_render_async": (C.c_int, [VP, C.c_uint32]),
    "paths_renderer_sync": (C.c_int, [VP]),
    "paths_renderer_combine": (C.c_int, [VP, C.c_int32]),
    "paths_renderer_combine_async": (C.c_int, [VP, C.c_int32]),
    "paths_renderer_attach_comm": (C.c_int, [VP, VP]),
    "paths_comm_unique_id": (C.c_int, [VP]),
    "paths_nccl_info": (C.c_char_p, []),
    "paths_group_create": (C.c_int, [VP, C.POINTER(Camera), C.POINTER(RenderOpts), C.c_int32, C.POINTER(C.c_int32), C.POINTER(VP)]),
    "paths_group_render": (C.c_int, [VP, C.c_uint32]),
    "paths_group_image_f32": (C.c_int, [VP, VP]),
    "paths_group_image_rgb24": (C.c_int, [VP, VP]),
    "paths_group_get_sums": (C.c_int, [VP, VP]),
    "paths_group_get_stats": (C.c_int, [VP, C.POINTER(Stats)]),
    "paths_group_renderer": (C.c_int, [VP, C.c_int32, C.POINTER(VP)]),
    "paths_group_set_camera": (C.c_int, [VP, VP, VP]),
    "paths_group_reset": (C.c_int, [VP]),
    "paths_group_destroy": (None, [VP]),
    "paths_renderer_image_f32": (C.c_int, [VP, VP]),
    "paths_renderer_image_rgb24": (C.c_int, [VP, VP]),
    "paths_renderer_get_sums": (C.c_int, [VP, VP]),
    "paths_renderer_add_sums": (C.c_int, [VP, VP]),
    "paths_renderer_device_sums": (C.c_int, [VP, C.POINTER(VP), C.POINTER(C.c_size_t), C.POINTER(C.c_int32)]),
    "paths_renderer_cuda_stream": (C.c_int, [VP, C.POINTER(VP)]),
    "paths_renderer_get_stats": (C.c_int, [VP, C.POINTER(Stats)]),
    "paths_renderer_set_camera": (C.c_int, [VP, VP, VP]),
    "paths_renderer_reset": (C.c_int, [VP]),
    "paths_renderer_render_preview": (C.c_int, [VP]),
    "paths_renderer_render_columns": (C.c_int, [VP, C.c_uint32]),
    "paths_renderer_get_progress": (C.c_int, [VP, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "paths_renderer_save_checkpoint": (C.c_int, [VP, C.c_char_p]),
    "paths_renderer_load_checkpoint": (C.c_int, [VP, C.c_char_p]),
    "paths_renderer_destroy": (None, [VP]),
    "paths_camera_rays": (C.c_int, [C.POINTER(Camera), C.c_int32, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                    VP, VP, VP, C.c_int32]),
    "paths_write_ppm": (C.c_int, [C.c_char_p, VP, C.c_uint32, C.c_uint32]),
    "paths_write_pfm": (C.c_int, [C.c_char_p, VP, C.c_uint32, C.c_uint32]),
}

_lib = None


def build(force=False):
    """Compiles the CUDA kernels (sm_100a) and the host runtime in-tree.  nvcc cross-compiles without a GPU."""
    args = ["make", "-C", str(_PKG), "-s", "-j8"]
    if force:
        args.append("-B")
    subprocess.check_call(args)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (there is no fallback path)")
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class PathsError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"paths_b200 status {status}: {message}")
        self.status = status


def check(status):
    if status != OK:
        raise PathsError(status, lib().paths_last_error().decode("utf-8", "replace"))
