/*
 * oracle.h -- C ABI of the CPU oracle for the paths render hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This library is a CPU (f64) restatement of the
 * reference renderer's algorithm (rynorris/paths, the files under src/).  It exists so the
 * CUDA product can be checked against it and so the CPU baseline can be timed.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load it.  Nothing under paths_b200/ links or calls it.
 *
 * Parity status: the reference cannot be compiled here (no rustc/cargo), so the
 * oracle is pinned against the reference's own unit-test vectors only
 * (bvh::test_morton_code, matrix::test_matrix_multiply, geom::switch_basis x4,
 * worker::pixel_grid_iter x4, sampling::test_sampler property tests x4).
 * Everything else (intersection, traversal, BSDFs, integrator, camera) is
 * "parity unpinned": the reference's tests hold no vectors for it; the source
 * is the only specification and each function below cites the lines it follows.
 */
#ifndef PATHS_ORACLE_H
#define PATHS_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- flattened scene description (what src/serde.rs deserialises) ---- */

typedef struct {
    uint32_t n_vertices;
    uint32_t n_faces;
    const double *vertices;       /* n_vertices x 3 (already f32-rounded by the loader, as tobj/ply-rs do) */
    const uint32_t *faces;        /* n_faces x 3 */
    const double *vertex_colours; /* n_vertices x 3 or NULL (src/ply.rs:59-71) */
} orc_model_desc;

enum { ORC_MAT_LAMBERTIAN = 0, ORC_MAT_MIRROR = 1, ORC_MAT_GLOSS = 2, ORC_MAT_COOK_TORRANCE = 3, ORC_MAT_FRESNEL = 4 };

typedef struct {
    int32_t type;        /* ORC_MAT_* (0..3 for a basic material) */
    int32_t albedo_mode; /* 0 = static rgb, 1 = per-vertex (src/material.rs:168-172) */
    double albedo[3];
    double reflectance;  /* Gloss r0 */
    double metalness;    /* Gloss */
    double roughness;    /* CookTorrance */
} orc_basic_material;

typedef struct {
    orc_basic_material m;        /* m.type == ORC_MAT_FRESNEL -> use diffuse/specular */
    double refractive_index;
    orc_basic_material diffuse;
    orc_basic_material specular;
} orc_material_desc;

typedef struct {
    int32_t shape;          /* 0 = Sphere, 1 = Mesh (one per loaded sub-model, src/serde.rs:112-141) */
    int32_t model;          /* index into models for Mesh */
    int32_t smooth_normals; /* src/serde.rs:243-252, default true */
    int32_t _pad;
    double center[3];
    double radius;
    double translation[3];
    double rot_pitch, rot_yaw, rot_roll; /* YAML values; the swapped-argument quirk of serde.rs:103 is applied by the oracle */
    double scale;
    orc_material_desc material;
} orc_object_desc;

typedef struct {
    int32_t geometry; /* 0 = Point, 1 = Sphere (src/serde.rs:213-229) */
    int32_t _pad;
    double v[3];      /* point vector or sphere centre */
    double radius;
    double colour[3];
    double intensity;
} orc_light_desc;

typedef struct {
    int32_t type; /* 0 = Flat, 1 = Gradient, 2 = Hdri (src/scene.rs:68-73) */
    uint32_t width, height;
    uint32_t _pad;
    double c0[3]; /* flat colour / overhead colour */
    double c1[3]; /* horizon colour */
    const float *texels; /* width*height*3 f32 RGB (image::hdr decode), row-major */
} orc_sky_desc;

typedef struct {
    uint32_t n_models;
    uint32_t n_objects;
    uint32_t n_lights;
    uint32_t _pad;
    const orc_model_desc *models;
    const orc_object_desc *objects;
    const orc_light_desc *lights;
    orc_sky_desc sky;
} orc_scene_desc;

typedef struct {
    uint32_t image_width, image_height;
    double location[3];
    double pitch, yaw, roll;
    double sensor_width, sensor_height, focal_length, focus_distance, aperture;
} orc_camera_desc;

typedef struct {
    uint64_t samples;      /* camera paths (what Renderer::num_rays_cast counts, src/renderer.rs:101) */
    uint64_t extend_rays;  /* find_intersection calls from trace.rs:18 */
    uint64_t shadow_rays;  /* find_intersection calls from trace.rs:61 */
    uint64_t box_tests;    /* ray_box_collide calls */
    uint64_t tri_tests;    /* TrianglePrimitive::intersect calls */
    uint64_t sphere_tests; /* SpherePrimitive::intersect calls */
    uint64_t segments;     /* loop iterations of trace_ray that found a hit or a miss */
    uint64_t bad_samples;  /* Colour::check() would have panicked */
} orc_counters;

typedef struct {
    int32_t type;      /* 0 sphere, 1 triangle */
    int32_t is_light;  /* entity kind */
    uint32_t entity;   /* object id / light id */
    uint32_t face;     /* triangle face index in its model */
    double v[9];       /* triangle vertices or (centre, radius, 0...) */
    double normal[3];
    double aabb_min[3], aabb_max[3];
} orc_prim_info;

typedef struct orc_scene orc_scene;

/* scene */
orc_scene *orc_scene_create(const orc_scene_desc *desc);
void orc_scene_destroy(orc_scene *s);
const char *orc_last_error(void);
uint32_t orc_scene_num_prims(const orc_scene *s);
uint32_t orc_scene_num_nodes(const orc_scene *s);
uint32_t orc_scene_bvh_depth(const orc_scene *s);
int orc_scene_prim_info(const orc_scene *s, uint32_t prim, orc_prim_info *out);

/* BVH::find_intersection over a batch of rays (src/bvh.rs:78-140). prim = 0xFFFFFFFF on miss. */
void orc_trace_closest(const orc_scene *s, const double *origins, const double *dirs, size_t n,
                       uint32_t *prim, double *t, double *bary /* n x 3 or NULL */, orc_counters *ctr /* or NULL */);
/* one primitive against one ray; returns 1 on hit */
int orc_intersect_prim(const orc_scene *s, uint32_t prim, const double o[3], const double d[3], double *t);

/* camera rays for sample index S of every pixel of the rectangle; o,d: n x 3; weight: n */
void orc_camera_rays(const orc_camera_desc *cam, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0,
                     uint32_t w, uint32_t h, double *o, double *d, double *weight);

/* full render: samples [s0, s1) of every pixel owned by (rank, world) under the shard mode
 * (mode 0: every pixel; mode 1: 64x64 tiles dealt round-robin, tile t -> rank t % world).
 * sums: W*H*3 f64 (+=), counts: W*H u32 (+=). */
void orc_render(const orc_scene *s, const orc_camera_desc *cam, uint64_t seed, uint32_t s0, uint32_t s1,
                int shard_mode, uint32_t rank, uint32_t world, int threads,
                double *sums, uint32_t *counts, orc_counters *ctr);

/* single path, for unit tests: returns colour*weight of sample S at pixel (x,y) */
void orc_trace_pixel(const orc_scene *s, const orc_camera_desc *cam, uint64_t seed, uint32_t x, uint32_t y,
                     uint32_t sample, double out_rgb[3]);

/* ---- unit-level entry points used by the KAT tests ---- */
uint32_t orc_permute(uint32_t i, uint32_t l, uint32_t p);
double orc_rand_float(uint32_t i, uint32_t p);
void orc_cmj(uint32_t s, uint32_t m, uint32_t n, uint32_t p, double out[2]);
void orc_cmj_disk(uint32_t s, uint32_t m, uint32_t n, uint32_t p, double out[2]);
uint64_t orc_morton_code(uint16_t num_bits, uint16_t x, uint16_t y, uint16_t z);
void orc_matmul3(const double a[9], const double b[9], double out[9]);
void orc_rotation(double yaw, double pitch, double roll, double out[9]);
void orc_switch_basis(const double v[3], const double normal[3], double out[3]);
void orc_to_bytes(const double *rgb, size_t n_pixels, uint8_t *out);
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void orc_pixel_grid(uint32_t x, uint32_t y, uint32_t w, uint32_t h, uint32_t *out_xy /* w*h*2 */);
void orc_sample_schedule(uint32_t sample, uint32_t out[4] /* m, n, s, pattern */);
void orc_sky(const orc_scene *s, const double dir[3], double out[3]);
void orc_material_brdf(const orc_material_desc *m, const double vec_out[3], const double vec_in[3],
                       const double normal[3], double out[3]);
/* u = {gloss choice, hemisphere u, hemisphere v}; returns is_specular */
int orc_material_sample(const orc_material_desc *m, const double vec_out[3], const double normal[3],
                        const double u[3], double dir[3], double *pdf, double brdf[3]);
void orc_light_sample(const orc_light_desc *l, const double from[3], double u, double v, double dir[3], double *inv_pdf);
uint32_t orc_ccrf(uint32_t x);

#ifdef __cplusplus
}
#endif
#endif
