//! `Renderer` with the reference's method names over the C ABI (UNBUILT here: no Rust toolchain in the development image).
//!
//! Drop-in plan for rynorris/paths: keep `main.rs`, `controller.rs`, `serde.rs`; replace `src/renderer.rs` + `src/worker.rs` by
//! `pub use paths_b200::Renderer;`.  `Controller` calls `fill_request_queue` / `drain_result_queue` once per UI tick
//! (`src/controller.rs:26-36`) and `render()` for the texture upload (`src/main.rs:114-125`); those keep their signatures.
use paths_b200_sys as sys;
use std::ffi::{CStr, CString};
use std::ptr;

#[derive(Debug)]
pub struct Error { pub status: i32, pub message: String }

fn check(status: i32) -> Result<(), Error> {
    if status == sys::PATHS_OK { return Ok(()); }
    let message = unsafe { CStr::from_ptr(sys::paths_last_error()) }.to_string_lossy().into_owned();
    Err(Error { status, message })
}

/// `Image {width, height, pixels}` of `src/camera.rs:6-10`: row-major linear-radiance means, index `x + y * width`, rgb triples.
pub struct Image { pub width: u32, pub height: u32, pub pixels: Vec<f32> }

pub struct Scene { raw: *mut sys::paths_scene, pub camera: sys::paths_camera }

impl Scene {
    /// `serde_yaml::from_reader` + `SceneDescription::{camera, scene}` (`src/main.rs:43-53`); width / height 0 keep the file's values.
    pub fn load(yaml_path: &str, width: u32, height: u32) -> Result<Scene, Error> {
        let path = CString::new(yaml_path).unwrap();
        let mut raw = ptr::null_mut();
        let mut camera: sys::paths_camera = unsafe { std::mem::zeroed() };
        check(unsafe { sys::paths_scene_load_yaml(path.as_ptr(), ptr::null(), width, height, &mut raw, &mut camera) })?;
        Ok(Scene { raw, camera })
    }
    /// `generate_stress_scene` (`src/stress.rs:7-66`), seeded.
    pub fn stress(num_spheres: u32, seed: u64) -> Result<Scene, Error> {
        let mut raw = ptr::null_mut();
        let mut camera: sys::paths_camera = unsafe { std::mem::zeroed() };
        check(unsafe { sys::paths_scene_create_stress(num_spheres, seed, &mut raw, &mut camera) })?;
        Ok(Scene { raw, camera })
    }
    /// `BVH::find_intersection` over a batch of rays (`src/bvh.rs:78-140`): (primitive id or PATHS_MISS, distance) per ray.
    pub fn find_intersection(&self, origins: &[[f64; 3]], dirs: &[[f64; 3]], precision: i32) -> Result<(Vec<u32>, Vec<f64>), Error> {
        assert_eq!(origins.len(), dirs.len());
        let n = origins.len();
        let (mut prim, mut t) = (vec![0u32; n], vec![0f64; n]);
        check(unsafe { sys::paths_trace_closest(self.raw, 0, origins.as_ptr() as *const f64, dirs.as_ptr() as *const f64, n,
                                                 prim.as_mut_ptr(), t.as_mut_ptr(), ptr::null_mut(), precision) })?;
        Ok((prim, t))
    }
    /// The f32 production closest hit with no device allocation (scratch in a static device array of the library): at most
    /// `PATHS_STATIC_MAX_RAYS` rays per call, and the scene must already be staged on the device (any earlier query or renderer).
    /// No counterpart in the reference.  The caller's output slices are filled in place: nothing is allocated on the host either.
    pub fn find_intersection_static(&self, origins: &[[f64; 3]], dirs: &[[f64; 3]], prim: &mut [u32], t: &mut [f64]) -> Result<(), Error> {
        assert!(origins.len() == dirs.len() && prim.len() == origins.len() && t.len() == origins.len());
        check(unsafe { sys::paths_trace_closest_static(self.raw, 0, origins.as_ptr() as *const f64, dirs.as_ptr() as *const f64, origins.len(),
                                                        prim.as_mut_ptr(), t.as_mut_ptr(), ptr::null_mut()) })
    }
}
impl Drop for Scene { fn drop(&mut self) { unsafe { sys::paths_scene_destroy(self.raw) } } }

/// `Renderer` (`src/renderer.rs:33-141`) on one GPU.  `num_workers` is accepted and ignored: the thread pool is a grid of persistent warps.
pub struct Renderer { raw: *mut sys::paths_renderer, width: u32, height: u32, samples_per_tick: u32 }

impl Renderer {
    pub fn new(scene: &Scene, _num_workers: usize) -> Result<Renderer, Error> {
        let mut opts: sys::paths_render_opts = unsafe { std::mem::zeroed() };
        unsafe { sys::paths_render_opts_default(&mut opts) };
        Renderer::with_opts(scene, &opts)
    }
    pub fn with_opts(scene: &Scene, opts: &sys::paths_render_opts) -> Result<Renderer, Error> {
        let mut raw = ptr::null_mut();
        check(unsafe { sys::paths_renderer_create(scene.raw, &scene.camera, opts, &mut raw) })?;
        Ok(Renderer { raw, width: scene.camera.image_width, height: scene.camera.image_height, samples_per_tick: 1 })
    }
    /// renderer.rs:79-91: the reference queues up to 200 column requests per tick; here one call advances every pixel.
    pub fn fill_request_queue(&mut self) -> Result<(), Error> { check(unsafe { sys::paths_renderer_render_async(self.raw, self.samples_per_tick) }) }
    /// renderer.rs:93-110
    pub fn drain_result_queue(&mut self) -> Result<(), Error> { check(unsafe { sys::paths_renderer_sync(self.raw) }) }
    pub fn set_samples_per_tick(&mut self, spp: u32) { self.samples_per_tick = spp.max(1); }
    /// renderer.rs:71-73
    pub fn render(&mut self) -> Result<Image, Error> {
        let mut pixels = vec![0f32; (self.width as usize) * (self.height as usize) * 3];
        check(unsafe { sys::paths_renderer_image_f32(self.raw, pixels.as_mut_ptr()) })?;
        Ok(Image { width: self.width, height: self.height, pixels })
    }
    /// `Colour::to_bytes` packed as RGB24 (colour.rs:62-74, main.rs:114-120): write straight into the SDL texture buffer.
    pub fn render_rgb24(&mut self, out: &mut [u8]) -> Result<(), Error> {
        assert!(out.len() >= (self.width as usize) * (self.height as usize) * 3);
        check(unsafe { sys::paths_renderer_image_rgb24(self.raw, out.as_mut_ptr()) })
    }
    /// renderer.rs:75-77 (counts camera samples)
    pub fn num_rays_cast(&mut self) -> Result<u64, Error> {
        let mut st = sys::paths_stats::default();
        check(unsafe { sys::paths_renderer_get_stats(self.raw, &mut st) })?;
        Ok(st.samples)
    }
    /// renderer.rs:112-120: `orientation` row-major 3x3; starts a new epoch
    pub fn set_camera(&mut self, location: [f64; 3], orientation: [f64; 9]) -> Result<(), Error> {
        check(unsafe { sys::paths_renderer_set_camera(self.raw, location.as_ptr(), orientation.as_ptr()) })
    }
    /// renderer.rs:122-128
    pub fn reset(&mut self) -> Result<(), Error> { check(unsafe { sys::paths_renderer_reset(self.raw) }) }
    /// renderer.rs:152-164
    pub fn request_preview(&mut self) -> Result<(), Error> { check(unsafe { sys::paths_renderer_render_preview(self.raw) }) }
    /// renderer.rs:166-192: the next `n` centre-out column requests
    pub fn render_columns(&mut self, n: u32) -> Result<(), Error> { check(unsafe { sys::paths_renderer_render_columns(self.raw, n) }) }
    /// renderer.rs:130-141
    pub fn shutdown(self) {}
}
impl Drop for Renderer { fn drop(&mut self) { unsafe { sys::paths_renderer_destroy(self.raw) } } }

/// One frame on several GPUs of this process; the partial estimators are combined on the first device inside the library over NCCL.
pub struct RendererGroup { raw: *mut sys::paths_group, width: u32, height: u32 }

impl RendererGroup {
    pub fn new(scene: &Scene, devices: &[i32], tiles: bool) -> Result<RendererGroup, Error> {
        let mut opts: sys::paths_render_opts = unsafe { std::mem::zeroed() };
        unsafe { sys::paths_render_opts_default(&mut opts) };
        opts.shard_mode = if tiles { sys::PATHS_SHARD_TILES } else { sys::PATHS_SHARD_SAMPLES };
        let mut raw = ptr::null_mut();
        check(unsafe { sys::paths_group_create(scene.raw, &scene.camera, &opts, devices.len() as i32, devices.as_ptr(), &mut raw) })?;
        Ok(RendererGroup { raw, width: scene.camera.image_width, height: scene.camera.image_height })
    }
    pub fn render_samples(&mut self, spp: u32) -> Result<(), Error> { check(unsafe { sys::paths_group_render(self.raw, spp) }) }
    pub fn render(&mut self) -> Result<Image, Error> {
        let mut pixels = vec![0f32; (self.width as usize) * (self.height as usize) * 3];
        check(unsafe { sys::paths_group_image_f32(self.raw, pixels.as_mut_ptr()) })?;
        Ok(Image { width: self.width, height: self.height, pixels })
    }
}
impl Drop for RendererGroup { fn drop(&mut self) { unsafe { sys::paths_group_destroy(self.raw) } } }
