"""Calls every C-ABI entry point with NULL handles / NULL outputs / out-of-range ids (run as a subprocess by test_host_cpu.py so that a
crash shows up as a failed test): each must return a paths_status error and set paths_last_error(), none may succeed or fault."""
import ctypes as C, sys
sys.path.insert(0, str(__import__('pathlib').Path(__file__).resolve().parent.parent))
import paths_b200 as pb
from paths_b200 import _capi
L = C.CDLL(str(_capi.LIB_PATH))   # raw handle: no argtypes, so None goes through as NULL
L.paths_last_error.restype = C.c_char_p
bad = []
def call(name, *a):
    f = getattr(L, name); f.restype = C.c_int
    rc = f(*a)
    msg = L.paths_last_error() or b''
    print(f"{name}{tuple(type(x).__name__ for x in a)} -> {rc} {msg[:70]!r}", flush=True)
    if rc == 0: bad.append(name)
    return rc
h = C.c_void_p(); cam = _capi.Camera(); info = _capi.SceneInfo(); opts = _capi.RenderOpts(); st = _capi.Stats()
u32 = C.c_uint32(); i32 = C.c_int32(); vp = C.c_void_p(); sz = C.c_size_t()
buf = (C.c_double * 16)()
call('paths_device_count', None)
call('paths_scene_create', None, C.byref(h)); call('paths_scene_create', C.byref(_capi.SceneDesc()), None)
call('paths_scene_load_yaml', None, None, 0, 0, C.byref(h), C.byref(cam))
call('paths_scene_load_yaml', b'scenes/teapot.yml', None, 0, 0, None, C.byref(cam))
call('paths_scene_create_stress', 10, 1, None, C.byref(cam))
L.paths_scene_destroy.restype = None; L.paths_scene_destroy(None)
call('paths_scene_build_bvh', None, 0)
call('paths_scene_check_bvh', None, 0, C.byref(u32))
call('paths_scene_get_info', None, C.byref(info))
call('paths_scene_prim_geometry', None, 0, buf, C.byref(i32))
call('paths_scene_prim_entity', None, 0, C.byref(i32), C.byref(u32), C.byref(u32))
call('paths_trace_closest', None, 0, buf, buf, C.c_size_t(1), C.byref(u32), buf, None, 64)
call('paths_trace_closest_static', None, 0, buf, buf, C.c_size_t(1), C.byref(u32), buf, None)
u8 = C.c_uint8()
call('paths_trace_occluded', None, 0, buf, buf, None, C.c_size_t(1), C.byref(u8), 32)
L.paths_render_opts_default.restype = None; L.paths_render_opts_default(None)
call('paths_renderer_create', None, C.byref(cam), C.byref(opts), C.byref(h))
for name in ('paths_renderer_sync', 'paths_renderer_reset', 'paths_renderer_render_preview'):
    call(name, None)
call('paths_renderer_render', None, 1); call('paths_renderer_render_async', None, 1); call('paths_renderer_render_columns', None, 1)
call('paths_renderer_image_f32', None, buf); call('paths_renderer_image_rgb24', None, buf)
call('paths_renderer_get_sums', None, buf); call('paths_renderer_add_sums', None, buf)
call('paths_renderer_device_sums', None, C.byref(vp), C.byref(sz), C.byref(i32))
call('paths_renderer_cuda_stream', None, C.byref(vp)); call('paths_renderer_get_stats', None, C.byref(st))
call('paths_renderer_set_camera', None, buf, buf); call('paths_renderer_get_progress', None, C.byref(u32), C.byref(u32))
call('paths_renderer_save_checkpoint', None, b'/tmp/x.ck'); call('paths_renderer_load_checkpoint', None, b'/tmp/x.ck')
L.paths_renderer_destroy.restype = None; L.paths_renderer_destroy(None)
call('paths_camera_rays', None, 0, C.c_uint64(1), 0, 0, 0, 1, 1, 64, buf, buf, buf)
call('paths_write_ppm', None, buf, 1, 1); call('paths_write_ppm', b'/tmp/x.ppm', None, 1, 1); call('paths_write_pfm', None, buf, 1, 1)
# a real scene with null outputs
s = C.c_void_p()
assert call.__globals__['L'].paths_scene_load_yaml(b'scenes/spheres_on_plane.yml', None, 0, 0, C.byref(s), C.byref(cam)) == 0
call('paths_scene_get_info', s, None)
call('paths_scene_prim_geometry', s, 10**6, buf, C.byref(i32))
call('paths_scene_prim_geometry', s, 0, None, C.byref(i32))
call('paths_scene_prim_entity', s, 10**6, C.byref(i32), C.byref(u32), C.byref(u32))
call('paths_scene_build_bvh', s, 99)
call('paths_trace_closest', s, 0, None, buf, C.c_size_t(1), C.byref(u32), buf, None, 64)
call('paths_trace_closest', s, 0, buf, buf, C.c_size_t(1), None, buf, None, 64)
call('paths_trace_closest', s, 0, buf, buf, C.c_size_t(1), C.byref(u32), buf, None, 17)
call('paths_trace_closest_static', s, 0, None, buf, C.c_size_t(1), C.byref(u32), buf, None)
call('paths_trace_closest_static', s, 0, buf, buf, C.c_size_t(1), C.byref(u32), None, None)
call('paths_trace_closest_static', s, 0, buf, buf, C.c_size_t(1), C.byref(u32), buf, None)  # no device here / scene not staged: refused either way
call('paths_trace_occluded', s, 0, buf, buf, None, C.c_size_t(1), None, 32)
call('paths_renderer_create', s, None, C.byref(opts), C.byref(h))
call('paths_renderer_create', s, C.byref(cam), C.byref(opts), None)
print("UNEXPECTED OK:", bad)
sys.exit(1 if bad else 0)
