/*
 * paths_b200.h -- C ABI of the B200-native render hot path of rynorris/paths.
 *
 * The reference is a monolithic Rust binary with no FFI seam (Cargo.toml has no [lib]); the seam the hot
 * path sits behind is the public API of `Renderer` (src/renderer.rs:33-141) plus the per-ray query
 * `BVH::find_intersection` (src/bvh.rs:78).  Every entry point below names the reference item it replaces.
 * INTEGRATION.md shows the Rust `extern "C"` block a maintainer would add on the reference side.
 *
 * Conventions
 *   - plain pointers and sizes only; all bulk buffers are caller-allocated HOST memory unless the name says
 *     `device`; the library copies what it keeps.
 *   - every function returns paths_status (0 = OK) and never unwinds; paths_last_error() returns a
 *     thread-local message for the last failure (the reference panics instead: renderer.rs:88,107-109).
 *   - handles are not thread-safe: one caller thread per renderer, like `&mut self` on Renderer.
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails with PATHS_ERR_CUDA.
 */
#ifndef PATHS_B200_H
#define PATHS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PATHS_OK = 0,
    PATHS_ERR_INVALID_ARG = 1,
    PATHS_ERR_IO = 2,
    PATHS_ERR_PARSE = 3,
    PATHS_ERR_CUDA = 4,
    PATHS_ERR_UNSUPPORTED = 5,
    PATHS_ERR_INTERNAL = 6,
    PATHS_ERR_NCCL = 7
} paths_status;

#define PATHS_MISS 0xFFFFFFFFu

/* ------------------------------------------------------------------ scene description
 * What src/serde.rs:65-74 deserialises (SceneDescription), flattened to C arrays. */

typedef struct {
    uint32_t n_vertices;
    uint32_t n_faces;
    const double *vertices;       /* n_vertices x 3; loaders deliver f32-rounded values like tobj / ply-rs */
    const uint32_t *faces;        /* n_faces x 3 (src/model.rs:105-114) */
    const double *vertex_colours; /* n_vertices x 3 in [0,1] or NULL (src/ply.rs:59-71) */
} paths_model_desc;

enum { PATHS_MAT_LAMBERTIAN = 0, PATHS_MAT_MIRROR = 1, PATHS_MAT_GLOSS = 2, PATHS_MAT_COOK_TORRANCE = 3, PATHS_MAT_FRESNEL = 4 };

typedef struct {
    int32_t type;        /* PATHS_MAT_* (src/material.rs:22-28) */
    int32_t albedo_mode; /* 0 = static rgb, 1 = per-vertex colour (src/material.rs:168-172) */
    double albedo[3];
    double reflectance; /* Gloss r0 (src/material.rs:282-300) */
    double metalness;   /* Gloss */
    double roughness;   /* CookTorrance m (src/material.rs:423-434) */
} paths_basic_material;

typedef struct {
    paths_basic_material m;    /* m.type == PATHS_MAT_FRESNEL selects diffuse/specular below */
    double refractive_index;   /* Fresnel (src/material.rs:373-387) */
    paths_basic_material diffuse;
    paths_basic_material specular;
} paths_material_desc;

typedef struct {
    int32_t shape;          /* 0 = Sphere, 1 = Mesh; one entry per loaded sub-model (src/serde.rs:91-142) */
    int32_t model;          /* Mesh: index into models */
    int32_t smooth_normals; /* src/serde.rs:243-252 (default true) */
    int32_t _pad;
    double center[3];       /* Sphere */
    double radius;
    double translation[3];  /* Mesh instance (src/geom.rs:98-122) */
    double rot_pitch, rot_yaw, rot_roll; /* YAML values, passed to Matrix3::rotation in the order of serde.rs:103 */
    double scale;
    paths_material_desc material;
} paths_object_desc;

typedef struct {
    int32_t geometry; /* 0 = Point, 1 = Sphere (src/serde.rs:213-229) */
    int32_t _pad;
    double v[3];      /* Point vector, or sphere centre */
    double radius;
    double colour[3];
    double intensity;
} paths_light_desc;

typedef struct {
    int32_t type; /* 0 = Flat, 1 = Gradient, 2 = Hdri (src/scene.rs:68-73) */
    uint32_t width, height;
    uint32_t _pad;
    double c0[3];        /* flat colour / overhead colour */
    double c1[3];        /* horizon colour */
    const float *texels; /* Hdri: width*height*3 f32 RGB, row-major (src/serde.rs:367-382) */
} paths_sky_desc;

typedef struct {
    uint32_t n_models;
    uint32_t n_objects;
    uint32_t n_lights;
    uint32_t _pad;
    const paths_model_desc *models;
    const paths_object_desc *objects;
    const paths_light_desc *lights;
    paths_sky_desc sky;
} paths_scene_desc;

/* CameraDescription (src/serde.rs:156-188) */
typedef struct {
    uint32_t image_width, image_height;
    double location[3];
    double pitch, yaw, roll;
    double sensor_width, sensor_height, focal_length, focus_distance, aperture;
} paths_camera;

typedef struct paths_scene paths_scene;
typedef struct paths_renderer paths_renderer;
typedef struct paths_group paths_group; /* one renderer per GPU of this process, rendering shards of one frame */

/* ------------------------------------------------------------------ options */

enum { PATHS_PRECISION_AUTO = 0, PATHS_PRECISION_F32 = 32, PATHS_PRECISION_MIXED = 48, PATHS_PRECISION_F64 = 64 };
/* f32 traversal kernels the per-ray seam can be pointed at (precision argument of paths_trace_closest / paths_trace_occluded);
 * PATHS_PRECISION_F32 itself means "whatever a render of this scene launches for bounce and visibility rays" */
enum { PATHS_TRACE_CAMERA = 33, PATHS_TRACE_TWO_WIDE = 34, PATHS_TRACE_PLAIN = 35, PATHS_TRACE_FOUR_WIDE = 36 };
enum { PATHS_SHARD_SAMPLES = 0, PATHS_SHARD_TILES = 1 };
enum { PATHS_BVH_FAST = 0, PATHS_BVH_REFERENCE = 1, PATHS_BVH_DEVICE = 2 };
/* Estimator extensions that have NO reference behaviour; all off by default so that output stays in parity.
 *   ENV_NEE        next-event estimation toward an Hdri sky through a CDF table (the reference only looks the map up on a miss,
 *                  src/scene.rs:95-111); a sky miss then counts only after a specular bounce, like light hits do (trace.rs:31-41)
 *   MICROFACET     importance-sample CookTorrance (Beckmann half vectors) and Fresnel combinations instead of ending the path
 *                  where Material::sample panics (src/material.rs:86)
 *   MICROFACET_COSINE  same materials sampled with a plain cosine lobe (cross-check of the estimator above)
 *   GGX            evaluate and sample CookTorrance with the GGX distribution instead of Beckmann (src/material.rs:436-447)
 *   NEE_ALL_LIGHTS weight the one-light NEE sample by the number of lights (the reference omits the factor, scene.rs:199-206) */
enum { PATHS_EXT_ENV_NEE = 1, PATHS_EXT_MICROFACET = 2, PATHS_EXT_MICROFACET_COSINE = 4, PATHS_EXT_GGX = 8, PATHS_EXT_NEE_ALL_LIGHTS = 16 };

typedef struct {
    int32_t device;        /* CUDA ordinal */
    int32_t precision;     /* PATHS_PRECISION_*: F32 = production kernels; F64 = parity instantiation (the reference's tree and visiting
                            * order); MIXED = f64 path state, shading and primitive tests over the production tree with f32 box tests --
                            * what AUTO picks for scenes whose coordinates are too large for f32 rays (beyond +-128 units) */
    int32_t shard_mode;    /* PATHS_SHARD_* */
    uint32_t rank, world;  /* this renderer's shard; world = 0 or 1 means unsharded */
    uint64_t seed;         /* key of the counter-based RNG (the reference's thread_rng is unseedable) */
    uint32_t max_bounces;  /* trace.rs:14 (10) */
    uint32_t samples_per_pass; /* camera samples per pixel kept in flight per wavefront pass; 0 = choose */
    int32_t time_kernels;  /* record CUDA events around every kernel (paths_stats.ms_*) */
    uint32_t extensions;   /* PATHS_EXT_* bit mask, 0 = the reference's estimator */
} paths_render_opts;

typedef struct {
    uint64_t samples;      /* camera paths = Renderer::num_rays_cast (src/renderer.rs:75-77,101) */
    uint64_t extend_rays;  /* closest-hit queries traced (trace.rs:18) */
    uint64_t shadow_rays;  /* visibility queries traced (trace.rs:61; provably-zero contributions are skipped) */
    uint64_t bad_samples;  /* Colour::check() would have panicked / unsampleable material (colour.rs:56-60, material.rs:86) */
    uint64_t kernel_launches;
    uint64_t passes;
    double ms_raygen, ms_extend, ms_shade, ms_shadow, ms_accumulate, ms_total; /* valid with time_kernels */
    uint64_t n_extend_launches, n_shadow_launches;
    double ms_device; /* device time of all render calls since reset: CUDA events on the launching stream, first to last kernel */
    /* Debug-assert builds only (make EXTRA=-DPATHS_DEBUG_ASSERTS; debug_asserts = 1): two-wide traversal stack overflows (a dropped
     * subtree = a possibly missed hit) and failed device assertions (queue positions inside their arrays) on this device since the
     * library was loaded.  Release builds bound the tree depth on the host instead (both builders refuse trees deeper than 60 levels for
     * the 64-entry stack) and report 0; the ray-level tests are run once under the debug-assert build (profiles/). */
    uint64_t stack_overflows, assert_failures;
    /* which kernels this renderer launches: extend / visibility kernel for bounce rays (0 plain batches, 3 two-wide lane refill,
     * 16 four-wide), camera segment (1 = raygen fused with a plain-batch closest hit), shade scheduling (1 static round-robin,
     * 0 dynamic fetch), pass contexts in flight */
    int32_t extend_kernel, camera_kernel, shade_kernel, pass_contexts;
    int32_t debug_asserts, _pad;
} paths_stats;

typedef struct {
    uint32_t n_prims;           /* flattened items: objects in order (mesh faces in face order), then area lights (scene.rs:143-170) */
    uint32_t n_triangles, n_spheres;
    uint32_t n_objects, n_lights;
    uint32_t ref_nodes, ref_depth;   /* reference-order binary AAC tree (bvh.rs:160-206); 0 until built */
    uint32_t fast_nodes, fast_depth; /* production wide tree */
    uint32_t hidden_prims;      /* prims whose own AABB is flat: never hit under bvh.rs:16's strict tmin < tmax */
    int32_t auto_precision;     /* what PATHS_PRECISION_AUTO resolves to for this scene */
    double bounds_min[3], bounds_max[3]; /* of prims with radius/extent < 1e5 */
    double build_ms_ref, build_ms_fast;
    /* production tree built on the GPU (PATHS_BVH_DEVICE); zeros until a device build has run */
    double build_ms_device;     /* CUDA-event time of the last device build */
    uint32_t bvh_builder;       /* builder of the production tree staged last: 0 host SAH, 1 device morton (LBVH) */
    uint32_t device_nodes2, device_nodes4, device_depth2, device_depth4;
    uint32_t giant_prims;       /* primitives chained above the root of the builder used last (the scenes' 1e6-radius plane spheres):
                                 * extent above 1024 x the geometric mean extent, at most 48; both builders apply the rule */
} paths_scene_info;

/* ------------------------------------------------------------------ scene (replaces SceneDescription::scene + Scene::new,
 * src/serde.rs:81-154, src/scene.rs:143-170; BVH build src/bvh.rs:160-206 runs on the host) */

const char *paths_last_error(void);
const char *paths_version(void);
paths_status paths_device_count(int32_t *count);

paths_status paths_scene_create(const paths_scene_desc *desc, paths_scene **out);
/* YAML -> scene + camera (src/main.rs:43-53).  asset_root may be NULL (paths resolved against the cwd, as the reference does,
 * then against the YAML's parent directory).  width/height = 0 keep the file's values. */
paths_status paths_scene_load_yaml(const char *yaml_path, const char *asset_root, uint32_t width, uint32_t height,
                                   paths_scene **out, paths_camera *camera);
/* generate_stress_scene (src/stress.rs:7-66; the scene src/main.rs:47-50 renders when no file is given, with 500 spheres): random
 * spheres with random Gloss / Lambertian / Mirror materials under a flat grey sky, no lights, and the reference's fixed camera.  The
 * reference draws from an unseedable thread_rng; here a SplitMix64 stream started at `seed` is consumed in the reference's order. */
paths_status paths_scene_create_stress(uint32_t num_spheres, uint64_t seed, paths_scene **out, paths_camera *camera);
void paths_scene_destroy(paths_scene *s);
/* construct_bvh_aac (src/bvh.rs:160-206) on the host, ahead of first use: which = PATHS_BVH_REFERENCE builds the reference-order
 * AAC tree the f64 kernels traverse, PATHS_BVH_FAST the production SAH tree.  Trees are otherwise built lazily. */
paths_status paths_scene_build_bvh(paths_scene *s, int32_t which);
/* which = PATHS_BVH_DEVICE selects the GPU builder for the production tree instead (morton codes, radix sort, Karras
 * hierarchy, bottom-up boxes with an SAH leaf decision -- milliseconds for ~1M triangles): it runs on each device at staging time;
 * scenes with fewer than 64 ordinary primitives keep the host builder.  PATHS_BVH_FAST switches back.  Either call drops the
 * production trees already staged, so renderers created afterwards use the new tree; existing renderers keep theirs.
 * paths_scene_check_bvh walks the production tree staged on `device` (staging it first if needed) on the host and counts
 * violations: a visible primitive missing from or repeated in the leaves, a primitive or child box sticking out of its parent's box. */
paths_status paths_scene_check_bvh(paths_scene *s, int32_t device, uint32_t *violations);
paths_status paths_scene_get_info(const paths_scene *s, paths_scene_info *info);
/* per-prim identity as the reference reports it: (is_light, entity id, face index) (bvh.rs T6, geom.rs:48-57) */
/* flattened geometry of one item after Mesh::primitives / TrianglePrimitive::transform (src/geom.rs:111-117,251-261):
 * triangle: a, b, c, n (12 doubles); sphere: centre, radius, then zeros.  *is_triangle may be NULL. */
paths_status paths_scene_prim_geometry(const paths_scene *s, uint32_t prim, double out[12], int32_t *is_triangle);
paths_status paths_scene_prim_entity(const paths_scene *s, uint32_t prim, int32_t *is_light, uint32_t *entity, uint32_t *face);

/* BVH::find_intersection over a batch of host rays (src/bvh.rs:78-140).  prim_id = PATHS_MISS and t = +inf on a miss.
 * precision 64: reference-order binary tree, f64, reference operation order (bit-exact ids and t);
 * precision 48: production tree, f32 box tests, the reference's f64 primitive tests (mixed mode), through the lane-refill kernel a
 *   mixed-precision render launches for its bounce rays (k_extend_refill_m, k_extend4_m above 64k primitives);
 * precision 32: the PRODUCTION closest-hit kernel: the rays are staged into a path-state buffer and traced by the very kernel, node
 *   layout and grid a render of this scene launches for its bounce rays (k_extend_refill over two-wide nodes up to 64k primitives,
 *   k_extend4s over four-wide nodes above); f32 boxes and watertight f32 triangles, f64 spheres;
 * PATHS_TRACE_CAMERA: the routine the fused camera kernel runs on primary rays (plain batches of 32 over two-wide nodes);
 * PATHS_TRACE_TWO_WIDE / _FOUR_WIDE / _PLAIN: one of the f32 kernels regardless of the scene's size.
 * bary (n x 3) may be NULL. */
paths_status paths_trace_closest(paths_scene *s, int32_t device, const double *origins, const double *dirs, size_t n,
                                 uint32_t *prim_id, double *t, double *bary, int32_t precision);
/* The shadow query of trace_ray (src/trace.rs:61-66: closest hit, then `entity != light`) over a batch of host rays, through the
 * visibility kernel a render launches (precision 32: k_shadow_refill / k_shadow4 -- any-hit with the nearest hit on the light's own
 * sphere as the search bound; 48: the same predicate with f64 primitives (k_shadow_refill_m / k_shadow4_m); 64: the reference's closest-hit-then-compare rule on its
 * own tree; PATHS_TRACE_* codes as above).  light[i] = index of the light ray i was sampled toward (hits on that light do not
 * occlude), or NULL: every hit occludes.  occluded[i] = 1 blocked, 0 clear (the kernel added the record's contribution to its path's
 * radiance exactly once). */
paths_status paths_trace_occluded(paths_scene *s, int32_t device, const double *origins, const double *dirs, const uint32_t *light,
                                  size_t n, uint8_t *occluded, int32_t precision);

/* paths_trace_closest (precision 32) for callers that may not allocate device memory.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same
 * rays in, same answers out (the production closest-hit kernel a render of this scene launches for its bounce rays; bary may be NULL),
 * but every byte of device scratch comes from one file-scope `static __device__` array inside the library -- PATHS_STATIC_SCRATCH_BYTES
 * (8.25 MiB) that every GPU which loads the library holds until exit -- and the call makes no cudaMalloc / cudaFree or any allocating
 * library call.  Limits and rules:
 *   - n <= PATHS_STATIC_MAX_RAYS (65536) per call, else PATHS_ERR_INVALID_ARG with the limit in paths_last_error(); split larger batches;
 *   - the scene must ALREADY be staged on `device` with its production tree (staging allocates): any earlier paths_trace_closest /
 *     paths_trace_occluded with precision 32 or 48, paths_scene_check_bvh or paths_renderer_create on that device does it; otherwise
 *     PATHS_ERR_UNSUPPORTED;
 *   - all calls share the one array, which keeps the last call's contents; each call rewrites what it reads.  Calls that overlap on one
 *     GPU would race on it: the library serialises the calls of a process with a lock and each call waits for the device before it
 *     returns (every process has its own copy of the array). */
#define PATHS_STATIC_MAX_RAYS 65536u
#define PATHS_STATIC_SCRATCH_BYTES (256u + PATHS_STATIC_MAX_RAYS * 132u) /* counters + 132 B per ray: staged rays, records, answers */
paths_status paths_trace_closest_static(paths_scene *s, int32_t device, const double *origins, const double *dirs, size_t n,
                                        uint32_t *prim_id, double *t, double *bary);

/* ------------------------------------------------------------------ renderer (replaces Renderer, src/renderer.rs:33-141) */

void paths_render_opts_default(paths_render_opts *o);
/* Renderer::new(camera, scene, num_workers) (renderer.rs:34-69) */
paths_status paths_renderer_create(paths_scene *s, const paths_camera *cam, const paths_render_opts *opts, paths_renderer **out);
/* fill_request_queue + drain_result_queue (renderer.rs:79-110): blocking; advances every owned pixel by spp samples
 * (sample shard: this rank's share of [done, done+spp) ). */
paths_status paths_renderer_render(paths_renderer *r, uint32_t spp);
/* same, without waiting for the device (pair with paths_renderer_sync) */
paths_status paths_renderer_render_async(paths_renderer *r, uint32_t spp);
paths_status paths_renderer_sync(paths_renderer *r);
/* Renderer::render() -> Image (renderer.rs:71-73, pixels.rs:53-79): row-major means, index x + y*W; zero-count pixels are 0 */
paths_status paths_renderer_image_f32(paths_renderer *r, float *rgb /* W*H*3 */);
/* Colour::to_bytes + RGB24 pack (colour.rs:30-36,62-74; main.rs:114-120), pitch 3W */
paths_status paths_renderer_image_rgb24(paths_renderer *r, uint8_t *rgb /* W*H*3 */);
/* Estimator state (pixels.rs:6-31): W*H x {sum r, sum g, sum b, count} as f64; read for checkpointing or host-side combine */
paths_status paths_renderer_get_sums(paths_renderer *r, double *sums4 /* W*H*4 */);
/* adds W*H*4 f64 partial sums into the estimator (combine of shards / resume) */
paths_status paths_renderer_add_sums(paths_renderer *r, const double *sums4);
/* device-resident estimator buffer for in-place collectives (NCCL reduce over NVLink): W*H*4 elements of f32 (elem_bytes 4) or f64 (8) */
paths_status paths_renderer_device_sums(paths_renderer *r, void **device_ptr, size_t *n_elems, int32_t *elem_bytes);
/* the CUDA stream (cudaStream_t) every kernel of this renderer is launched on, so a caller can order collectives and
 * CUDA events against the render without a host synchronisation */
paths_status paths_renderer_cuda_stream(paths_renderer *r, void **stream);
/* num_rays_cast (renderer.rs:75-77) and friends */
paths_status paths_renderer_get_stats(paths_renderer *r, paths_stats *st);
/* set_camera(location, orientation) (renderer.rs:112-120): new epoch; rot is row-major 3x3 */
paths_status paths_renderer_set_camera(paths_renderer *r, const double location[3], const double rot[9]);
/* reset() (renderer.rs:122-128) */
paths_status paths_renderer_reset(paths_renderer *r);

/* ---- several GPUs, one frame.  In the reference the merge of partial results is part of Renderer itself
 * (drain_result_queue -> Estimator::update_pixel, src/renderer.rs:93-105); here it is paths_renderer_combine: the partial estimators
 * of all ranks of a frame are combined on `root` over NCCL (NVLink / NVSwitch), enqueued on the renderer's stream behind its passes.
 *   sample shards (PATHS_SHARD_SAMPLES): ncclReduce(sum) of the {sum r, sum g, sum b, count} buffers -- equal to one GPU up to the
 *     order of the f32 additions;
 *   tile shards (PATHS_SHARD_TILES): ranks own disjoint 32x32 tiles, so each packs its pixels and ncclSend's them, the root receives all
 *     of them in one NCCL group and scatters them -- bit-identical to one GPU, 1/world of the bytes of a full-frame reduce.
 * After a combine the root's image / estimator reads (image_f32, image_rgb24, get_sums) show the combined frame until the next sample is
 * rendered; each rank's own estimator is left untouched, so render -> combine -> render -> combine accumulates correctly.
 * NCCL is bound at run time (the libnccl.so.2 already loaded into the process, else the one on the loader path): PATHS_ERR_NCCL if none.
 *
 * One process per GPU (MPI / torchrun style): rank 0 calls paths_comm_unique_id and hands the 128 bytes to every rank out of band; every
 * rank then calls paths_renderer_attach_comm (collective: returns when all `world` ranks of paths_render_opts have joined). */
/* "NCCL 2.28.9 from /path/libnccl.so.2" once bound, else why it cannot be */
const char *paths_nccl_info(void);
#define PATHS_COMM_ID_BYTES 128
paths_status paths_comm_unique_id(uint8_t id[PATHS_COMM_ID_BYTES]);
paths_status paths_renderer_attach_comm(paths_renderer *r, const uint8_t id[PATHS_COMM_ID_BYTES]);
paths_status paths_renderer_combine(paths_renderer *r, int32_t root);       /* blocking */
paths_status paths_renderer_combine_async(paths_renderer *r, int32_t root); /* pair with paths_renderer_sync */
/* One process, several GPUs: the library owns one renderer, stream and NCCL rank per device (ncclCommInitAll).  opts->device, rank and
 * world are ignored (rank i renders on device_ids[i], world = n_devices); device_ids = NULL means devices 0 .. n_devices-1.
 * paths_group_render advances the frame by spp samples per pixel in total (sample shards: spp / n each; tile shards: all spp on every
 * rank's own tiles) and combines on device_ids[0]; the image calls return the combined frame. */
paths_status paths_group_create(paths_scene *s, const paths_camera *cam, const paths_render_opts *opts, int32_t n_devices,
                                const int32_t *device_ids, paths_group **out);
paths_status paths_group_render(paths_group *g, uint32_t spp);
paths_status paths_group_image_f32(paths_group *g, float *rgb /* W*H*3 */);
paths_status paths_group_image_rgb24(paths_group *g, uint8_t *rgb /* W*H*3 */);
paths_status paths_group_get_sums(paths_group *g, double *sums4 /* W*H*4 */);
/* counters summed over the devices; ms_device = the slowest device's */
paths_status paths_group_get_stats(paths_group *g, paths_stats *st);
/* the renderer of rank i, borrowed (per-device stats, stream, estimator); do not destroy it */
paths_status paths_group_renderer(paths_group *g, int32_t i, paths_renderer **out);
paths_status paths_group_set_camera(paths_group *g, const double location[3], const double rot[9]);
paths_status paths_group_reset(paths_group *g);
void paths_group_destroy(paths_group *g);

/* ---- the interactive front half of Renderer (src/renderer.rs:143-192), single GPU.
 * render_preview = request_preview (renderer.rs:152-164): one sample for every 6th pixel in x and y.  The image calls show a pixel
 * that has no samples yet through that grid, like Estimator::render (src/pixels.rs:53-79).
 * render_columns = the next n_requests of next_request (renderer.rs:166-192): whole pixel columns from the centre of the image
 * outwards, 1 sample per pixel during the first sweep of an epoch ((1,1) pattern), 25 per sweep afterwards ((5,5) pattern).  A sweep
 * of all columns advances every pixel exactly like paths_renderer_render(1) resp. (25) -- the estimator is bit-identical.
 * paths_renderer_render refuses to run in the middle of a sweep.  get_progress: samples per pixel completed by whole-frame calls and
 * finished sweeps, and the number of columns already rendered in the sweep in progress. */
paths_status paths_renderer_render_preview(paths_renderer *r);
paths_status paths_renderer_render_columns(paths_renderer *r, uint32_t n_requests);
paths_status paths_renderer_get_progress(paths_renderer *r, uint32_t *samples_per_pixel, uint32_t *sweep_position);
/* Checkpoint / resume (the reference keeps its Estimator in memory only and loses it on exit): the per-pixel sums and counts plus
 * the position in the sample schedule.  A renderer that loads a checkpoint (same image size and seed) and renders k more samples
 * holds the same estimator as one that rendered without interruption. */
paths_status paths_renderer_save_checkpoint(paths_renderer *r, const char *path);
paths_status paths_renderer_load_checkpoint(paths_renderer *r, const char *path);
/* shutdown() (renderer.rs:130-141) */
void paths_renderer_destroy(paths_renderer *r);

/* Camera::get_ray_for_pixel for sample index S of every pixel of a rectangle, on the device (camera.rs:47-94, worker.rs:58-87).
 * o, d: n x 3, weight: n (host). */
paths_status paths_camera_rays(const paths_camera *cam, int32_t device, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0,
                               uint32_t w, uint32_t h, double *o, double *d, double *weight, int32_t precision);

/* image writers for the headless sink (the reference only has an SDL window, main.rs:60-83) */
paths_status paths_write_ppm(const char *path, const uint8_t *rgb, uint32_t w, uint32_t h);
paths_status paths_write_pfm(const char *path, const float *rgb, uint32_t w, uint32_t h);

#ifdef __cplusplus
}
#endif
#endif
