//! Raw declarations of `include/paths_b200.h`.  UNBUILT in the image this repository is developed in (no rustc / cargo there).
#![allow(non_camel_case_types)]
use std::os::raw::c_char;

pub const PATHS_OK: i32 = 0;
pub const PATHS_ERR_INVALID_ARG: i32 = 1;
pub const PATHS_ERR_IO: i32 = 2;
pub const PATHS_ERR_PARSE: i32 = 3;
pub const PATHS_ERR_CUDA: i32 = 4;
pub const PATHS_ERR_UNSUPPORTED: i32 = 5;
pub const PATHS_ERR_INTERNAL: i32 = 6;
pub const PATHS_ERR_NCCL: i32 = 7;
pub const PATHS_MISS: u32 = 0xFFFF_FFFF;
pub const PATHS_PRECISION_AUTO: i32 = 0;
pub const PATHS_PRECISION_F32: i32 = 32;
pub const PATHS_PRECISION_MIXED: i32 = 48;
pub const PATHS_PRECISION_F64: i32 = 64;
pub const PATHS_STATIC_MAX_RAYS: usize = 65536;
pub const PATHS_TRACE_CAMERA: i32 = 33;
pub const PATHS_TRACE_TWO_WIDE: i32 = 34;
pub const PATHS_TRACE_PLAIN: i32 = 35;
pub const PATHS_TRACE_FOUR_WIDE: i32 = 36;
pub const PATHS_SHARD_SAMPLES: i32 = 0;
pub const PATHS_SHARD_TILES: i32 = 1;
pub const PATHS_BVH_FAST: i32 = 0;
pub const PATHS_BVH_REFERENCE: i32 = 1;
pub const PATHS_BVH_DEVICE: i32 = 2;
pub const PATHS_COMM_ID_BYTES: usize = 128;

#[repr(C)] pub struct paths_scene { _private: [u8; 0] }
#[repr(C)] pub struct paths_renderer { _private: [u8; 0] }
#[repr(C)] pub struct paths_group { _private: [u8; 0] }

#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_model_desc {
    pub n_vertices: u32, pub n_faces: u32,
    pub vertices: *const f64, pub faces: *const u32, pub vertex_colours: *const f64,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_basic_material {
    pub type_: i32, pub albedo_mode: i32, pub albedo: [f64; 3], pub reflectance: f64, pub metalness: f64, pub roughness: f64,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_material_desc {
    pub m: paths_basic_material, pub refractive_index: f64, pub diffuse: paths_basic_material, pub specular: paths_basic_material,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_object_desc {
    pub shape: i32, pub model: i32, pub smooth_normals: i32, pub _pad: i32,
    pub center: [f64; 3], pub radius: f64, pub translation: [f64; 3],
    pub rot_pitch: f64, pub rot_yaw: f64, pub rot_roll: f64, pub scale: f64,
    pub material: paths_material_desc,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_light_desc {
    pub geometry: i32, pub _pad: i32, pub v: [f64; 3], pub radius: f64, pub colour: [f64; 3], pub intensity: f64,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_sky_desc {
    pub type_: i32, pub width: u32, pub height: u32, pub _pad: u32, pub c0: [f64; 3], pub c1: [f64; 3], pub texels: *const f32,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_scene_desc {
    pub n_models: u32, pub n_objects: u32, pub n_lights: u32, pub _pad: u32,
    pub models: *const paths_model_desc, pub objects: *const paths_object_desc, pub lights: *const paths_light_desc,
    pub sky: paths_sky_desc,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_camera {            // CameraDescription, src/serde.rs:156-188
    pub image_width: u32, pub image_height: u32,
    pub location: [f64; 3],
    pub pitch: f64, pub yaw: f64, pub roll: f64,
    pub sensor_width: f64, pub sensor_height: f64, pub focal_length: f64, pub focus_distance: f64, pub aperture: f64,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_render_opts {
    pub device: i32, pub precision: i32, pub shard_mode: i32, pub rank: u32, pub world: u32,
    pub seed: u64, pub max_bounces: u32, pub samples_per_pass: u32, pub time_kernels: i32, pub extensions: u32,
}
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct paths_stats {
    pub samples: u64, pub extend_rays: u64, pub shadow_rays: u64, pub bad_samples: u64, pub kernel_launches: u64, pub passes: u64,
    pub ms_raygen: f64, pub ms_extend: f64, pub ms_shade: f64, pub ms_shadow: f64, pub ms_accumulate: f64, pub ms_total: f64,
    pub n_extend_launches: u64, pub n_shadow_launches: u64, pub ms_device: f64,
    pub stack_overflows: u64, pub assert_failures: u64,
    pub extend_kernel: i32, pub camera_kernel: i32, pub shade_kernel: i32, pub pass_contexts: i32, pub debug_asserts: i32, pub _pad: i32,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct paths_scene_info {
    pub n_prims: u32, pub n_triangles: u32, pub n_spheres: u32, pub n_objects: u32, pub n_lights: u32,
    pub ref_nodes: u32, pub ref_depth: u32, pub fast_nodes: u32, pub fast_depth: u32, pub hidden_prims: u32,
    pub auto_precision: i32,
    pub bounds_min: [f64; 3], pub bounds_max: [f64; 3],
    pub build_ms_ref: f64, pub build_ms_fast: f64, pub build_ms_device: f64,
    pub bvh_builder: u32, pub device_nodes2: u32, pub device_nodes4: u32, pub device_depth2: u32, pub device_depth4: u32, pub giant_prims: u32,
}

extern "C" {
    pub fn paths_last_error() -> *const c_char;
    pub fn paths_version() -> *const c_char;
    pub fn paths_nccl_info() -> *const c_char;
    pub fn paths_device_count(count: *mut i32) -> i32;
    // scene
    pub fn paths_scene_create(desc: *const paths_scene_desc, out: *mut *mut paths_scene) -> i32;
    pub fn paths_scene_load_yaml(yaml_path: *const c_char, asset_root: *const c_char, width: u32, height: u32,
                                 out: *mut *mut paths_scene, camera: *mut paths_camera) -> i32;
    pub fn paths_scene_create_stress(num_spheres: u32, seed: u64, out: *mut *mut paths_scene, camera: *mut paths_camera) -> i32;
    pub fn paths_scene_destroy(s: *mut paths_scene);
    pub fn paths_scene_build_bvh(s: *mut paths_scene, which: i32) -> i32;
    pub fn paths_scene_check_bvh(s: *mut paths_scene, device: i32, violations: *mut u32) -> i32;
    pub fn paths_scene_get_info(s: *const paths_scene, info: *mut paths_scene_info) -> i32;
    pub fn paths_scene_prim_geometry(s: *const paths_scene, prim: u32, out: *mut f64, is_triangle: *mut i32) -> i32;
    pub fn paths_scene_prim_entity(s: *const paths_scene, prim: u32, is_light: *mut i32, entity: *mut u32, face: *mut u32) -> i32;
    // per-ray seams (BVH::find_intersection, src/bvh.rs:78-140; the shadow query of trace_ray, src/trace.rs:61-66)
    pub fn paths_trace_closest(s: *mut paths_scene, device: i32, origins: *const f64, dirs: *const f64, n: usize,
                               prim_id: *mut u32, t: *mut f64, bary: *mut f64, precision: i32) -> i32;
    // no reference counterpart: paths_trace_closest (f32) with its scratch in a static device array, at most PATHS_STATIC_MAX_RAYS rays
    pub fn paths_trace_closest_static(s: *mut paths_scene, device: i32, origins: *const f64, dirs: *const f64, n: usize,
                                      prim_id: *mut u32, t: *mut f64, bary: *mut f64) -> i32;
    pub fn paths_trace_occluded(s: *mut paths_scene, device: i32, origins: *const f64, dirs: *const f64, light: *const u32, n: usize,
                                occluded: *mut u8, precision: i32) -> i32;
    // renderer (Renderer, src/renderer.rs:33-141)
    pub fn paths_render_opts_default(o: *mut paths_render_opts);
    pub fn paths_renderer_create(s: *mut paths_scene, cam: *const paths_camera, opts: *const paths_render_opts,
                                 out: *mut *mut paths_renderer) -> i32;
    pub fn paths_renderer_render(r: *mut paths_renderer, spp: u32) -> i32;
    pub fn paths_renderer_render_async(r: *mut paths_renderer, spp: u32) -> i32;
    pub fn paths_renderer_sync(r: *mut paths_renderer) -> i32;
    pub fn paths_renderer_image_f32(r: *mut paths_renderer, rgb: *mut f32) -> i32;
    pub fn paths_renderer_image_rgb24(r: *mut paths_renderer, rgb: *mut u8) -> i32;
    pub fn paths_renderer_get_sums(r: *mut paths_renderer, sums4: *mut f64) -> i32;
    pub fn paths_renderer_add_sums(r: *mut paths_renderer, sums4: *const f64) -> i32;
    pub fn paths_renderer_device_sums(r: *mut paths_renderer, device_ptr: *mut *mut std::ffi::c_void, n_elems: *mut usize, elem_bytes: *mut i32) -> i32;
    pub fn paths_renderer_cuda_stream(r: *mut paths_renderer, stream: *mut *mut std::ffi::c_void) -> i32;
    pub fn paths_renderer_get_stats(r: *mut paths_renderer, st: *mut paths_stats) -> i32;
    pub fn paths_renderer_set_camera(r: *mut paths_renderer, location: *const f64, rot: *const f64) -> i32;
    pub fn paths_renderer_reset(r: *mut paths_renderer) -> i32;
    pub fn paths_renderer_destroy(r: *mut paths_renderer);
    // several GPUs, one frame: the combine lives inside the library (src/renderer.rs:93-105)
    pub fn paths_comm_unique_id(id: *mut u8) -> i32;
    pub fn paths_renderer_attach_comm(r: *mut paths_renderer, id: *const u8) -> i32;
    pub fn paths_renderer_combine(r: *mut paths_renderer, root: i32) -> i32;
    pub fn paths_renderer_combine_async(r: *mut paths_renderer, root: i32) -> i32;
    pub fn paths_group_create(s: *mut paths_scene, cam: *const paths_camera, opts: *const paths_render_opts, n_devices: i32,
                              device_ids: *const i32, out: *mut *mut paths_group) -> i32;
    pub fn paths_group_render(g: *mut paths_group, spp: u32) -> i32;
    pub fn paths_group_image_f32(g: *mut paths_group, rgb: *mut f32) -> i32;
    pub fn paths_group_image_rgb24(g: *mut paths_group, rgb: *mut u8) -> i32;
    pub fn paths_group_get_sums(g: *mut paths_group, sums4: *mut f64) -> i32;
    pub fn paths_group_get_stats(g: *mut paths_group, st: *mut paths_stats) -> i32;
    pub fn paths_group_renderer(g: *mut paths_group, i: i32, out: *mut *mut paths_renderer) -> i32;
    pub fn paths_group_set_camera(g: *mut paths_group, location: *const f64, rot: *const f64) -> i32;
    pub fn paths_group_reset(g: *mut paths_group) -> i32;
    pub fn paths_group_destroy(g: *mut paths_group);
    // the interactive front half (src/renderer.rs:143-192), checkpoints
    pub fn paths_renderer_render_preview(r: *mut paths_renderer) -> i32;
    pub fn paths_renderer_render_columns(r: *mut paths_renderer, n_requests: u32) -> i32;
    pub fn paths_renderer_get_progress(r: *mut paths_renderer, samples_per_pixel: *mut u32, sweep_position: *mut u32) -> i32;
    pub fn paths_renderer_save_checkpoint(r: *mut paths_renderer, path: *const c_char) -> i32;
    pub fn paths_renderer_load_checkpoint(r: *mut paths_renderer, path: *const c_char) -> i32;
    // camera rays, image writers
    pub fn paths_camera_rays(cam: *const paths_camera, device: i32, seed: u64, sample: u32, x0: u32, y0: u32, w: u32, h: u32,
                             o: *mut f64, d: *mut f64, weight: *mut f64, precision: i32) -> i32;
    pub fn paths_write_ppm(path: *const c_char, rgb: *const u8, w: u32, h: u32) -> i32;
    pub fn paths_write_pfm(path: *const c_char, rgb: *const f32, w: u32, h: u32) -> i32;
}
