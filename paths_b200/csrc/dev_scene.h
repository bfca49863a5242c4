// dev_scene.h -- layout of the scene in HBM and of the wavefront path state (see DESIGN.md "Data layout").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace paths {

// prim_meta[prim] (prim = index in the reference's flattened item list, src/scene.rs:143-170)
//   x: bit0 = triangle, bit1 = entity is a light;  y: entity id;  z: face index in its model;  w: unused
#define PATHS_PM_TRI 1u
#define PATHS_PM_LIGHT 2u
#define PATHS_LIGHT_ENV 0xFFFFFFFEu /* shadow-record light id of an environment NEE sample */

// S = storage type of the real-valued fields: double for the parity / mixed instantiations, float for the production f32 kernels
// (the same values rounded once on the host instead of once per shaded hit: no 8-byte loads, no f64 -> f32 conversions in k_shade)
template <typename S> struct DevObjectT { // src/scene.rs:35-40 + src/geom.rs:98-122 + resolved material record
    int32_t shape, smooth;
    uint32_t vert_base, face_base;
    S rot[9];
    int32_t mat_type, vertex_albedo;
    S albedo[3];
    S r0, metalness, roughness;
    // Fresnel combination, brdf() only
    int32_t d_type, s_type;
    S d_albedo[3], s_albedo[3];
    S d_r0, d_metal, d_rough, s_r0, s_metal, s_rough, fresnel_r0;
};
typedef DevObjectT<double> DevObject;
typedef DevObjectT<float> DevObjectF;

struct DevLight { // src/scene.rs:42-66
    int32_t geometry;
    uint32_t prim;
    double v[3];
    double radius;
    double colour[3];
    double intensity;
};

struct DevRefNode { // reference-order binary tree, child boxes in the parent (src/bvh.rs:23-70)
    double lbox[6];
    double rbox[6];
    int32_t left, right; // >= 0 inner; < 0 leaf: ~prim
    int32_t pad[2];
};

struct DevScene {
    uint32_t n_prims, n_lights, n_objects;
    // ---- prim-id indexed tables
    const uint4 *prim_meta;
    const double *prim_geom; // 12 doubles per prim. triangle: a, b, c, n ; sphere: centre, radius
    // ---- reference tree (f64 parity kernels)
    const DevRefNode *ref_nodes;
    int32_t ref_root;
    double ref_root_box[6];
    // ---- production tree (f32 kernels): 4 float4 per node, 3 float4 per leaf slot
    const float4 *fast_nodes;
    const float4 *fast_leaves;
    int32_t fast_root;
    // ---- the same tree, four-wide: 8 float4 (128 B) per node = {lo.x, hi.x, lo.y, hi.y, lo.z, hi.z}[4 children], child codes, pad
    const float4 *fast4_nodes;
    int32_t fast4_root;
    uint32_t fast4_depth;
    // ---- shading tables
    const DevObject *objects;
    const DevLight *lights;
    const double *vertex_normals; // 3 per vertex, concatenated over models
    const double *vertex_colours; // 3 per vertex
    const uint32_t *faces;        // 3 per face
    // ---- f32 mirrors read by the production shade kernel (null unless the production tree is staged)
    const float4 *prim_shade;     // 2 per prim: {flags, entity, face, -} as bits | triangle: rotated face normal, sphere: centre
    const DevObjectF *objects32;
    const float4 *vertex_normals32; // xyz per vertex
    const float4 *vertex_colours32;
    int32_t sky_type;
    double sky_c0[3], sky_c1[3];
    uint32_t sky_w, sky_h;
    const float *sky_texels;
    const float *env_marginal;    // [sky_h + 1]
    const float *env_conditional; // [sky_h][sky_w + 1]
};

struct DevCamera { // src/camera.rs:12-27 after CameraDescription::to_camera (src/serde.rs:172-188)
    uint32_t width, height;
    double location[3];
    double rot[9];
    double sensor_width, sensor_height, focal_length, distance_from_lens, aperture;
};

// One wavefront pass keeps n_slots camera paths in flight.  Slot i carries sample (s_base + i / n_local) of local pixel
// (i % n_local).  The records a bounce reads and writes are DENSE BY QUEUE POSITION, not by slot: the shade kernel of bounce b writes
// the ray and throughput of every surviving path at the position it takes in bounce b + 1's queue, so that the closest-hit kernel and
// the next shade kernel stream contiguous lines (round 1 kept records in their slot and compacted indices only: after the first bounce
// every access dragged half-used sectors through DRAM -- 84 B per ray for 52 useful in the closest-hit kernel, more bytes per entry in
// shade bounce 1 than in bounce 0 -- and every ray fetch was a dependent load behind its queue entry).  Only the radiance record, which
// few bounces touch, stays indexed by slot.  Buffers alternate with the bounce parity p = b & 1.
template <typename R4> struct PathState {
    R4 *ray[2]; // ray[p][2 * pos] = {origin xyz, slot bits}, ray[p][2 * pos + 1] = {direction xyz, flags bits: loops | last_bounce_specular << 8}
    R4 *thr[2]; // thr[p][pos] = throughput rgb (bounce 0: implicitly (1, 1, 1), never stored)
    R4 *hit;    // hit[pos] = prim bits, t, bx, by of the bounce being traced
    R4 *col;    // col[slot] = radiance rgb gathered so far, w = camera weight (src/camera.rs:91)
    // shadow-ray records, dense by shadow queue position
    R4 *sh_o;  // origin xyz, w = light index bits
    R4 *sh_d;  // direction xyz, w = slot bits
    R4 *sh_c;  // contribution rgb (already multiplied by throughput)
};

#define PATHS_NWORK 4
struct PassControl {
    uint32_t *qcount;    // [max_bounces + 2] live paths entering bounce b
    uint32_t *scount;    // [max_bounces + 2] shadow rays emitted at bounce b
    uint32_t *ecount;    // [max_bounces + 2] environment NEE records emitted at bounce b (extension; second half of the shadow arrays)
    uint32_t *work;      // [PATHS_NWORK * (max_bounces + 2)] dynamic-fetch cursors, one per launch: extend, shade, shadow, env shadow
    unsigned long long *totals; // [0] extend rays [1] shadow rays [2] bad samples [3] samples
};

struct ShardInfo {
    uint32_t n_local;         // pixels owned by this rank
    const uint32_t *pixels;   // local -> global pixel index, or NULL for identity
    uint32_t n_local_magic;   // floor(2^32 / n_local): slot / n_local without an integer division (see shard_split)
    // Whole-frame passes (pixels == NULL) number their local pixels in 8 x 4 TILE order when the image size allows it (width % 8 == 0,
    // height % 4 == 0; tiles_x = width / 8, else 0 = row-major): 32 consecutive slots are then an 8 x 4 block of pixels instead of a
    // 32-pixel line, so the rays a warp generates, shades and re-emits start closer together and agree on more of the tree.  Results do
    // not depend on the numbering (every pixel's samples are a function of (seed, pixel, sample) only).
    uint32_t tiles_x, tiles_x_magic; // tiles_x_magic = floor(2^32 / tiles_x) + 1
    // Inside a tile the 32 pixels are in Morton order (x0 y0 x1 y1 x2), so the first 4 / 8 / 16 of them are a 2x2 / 4x2 / 4x4 block.
    // SAMPLE GROUPS: with sg_log2 = g > 0 a batch of 32 consecutive slots holds 2^g consecutive samples of 32 >> g consecutive local
    // pixels (g = 2: a 4 x 2 pixel block x 4 samples) instead of one sample of 32 pixels: the rays of a warp are then closer still.
    // Needs n_local % 32 == 0 and the pass's sample count % 2^g == 0 (else g = 0).  n_blocks = n_local >> (5 - g).
    uint32_t sg_log2, n_blocks, n_blocks_magic; // n_blocks_magic = floor(2^32 / n_blocks)
#if defined(__CUDACC__)
    __device__ __forceinline__ uint32_t pixel_of(uint32_t lp) const {
        if (pixels) return pixels[lp];
        if (tiles_x == 0) return lp;
        const uint32_t tile = lp >> 5, in = lp & 31u;
        uint32_t ty = __umulhi(tile, tiles_x_magic); // tile / tiles_x: exact for tile < 2^32 / tiles_x with this magic (checked on the host)
        uint32_t tx = tile - ty * tiles_x;
        const uint32_t x = (in & 1u) | ((in >> 1) & 2u) | ((in >> 2) & 4u), y = ((in >> 1) & 1u) | ((in >> 2) & 2u);
        return ((ty << 2) + y) * (tiles_x << 3) + (tx << 3) + x;
    }
    // slot -> (local pixel, sample within the pass); the estimate __umulhi(., magic) is at most one too small
    __device__ __forceinline__ void split(uint32_t slot, uint32_t &lp, uint32_t &k) const {
        if (sg_log2 == 0) {
            k = __umulhi(slot, n_local_magic);
            lp = slot - k * n_local;
            if (lp >= n_local) {
                lp -= n_local;
                k += 1;
            }
            return;
        }
        const uint32_t batch = slot >> 5, j = slot & 31u, pl = 5u - sg_log2;
        uint32_t sg = __umulhi(batch, n_blocks_magic);
        uint32_t pb = batch - sg * n_blocks;
        if (pb >= n_blocks) {
            pb -= n_blocks;
            sg += 1;
        }
        lp = (pb << pl) + (j & ((1u << pl) - 1u));
        k = (sg << sg_log2) + (j >> pl);
    }
    // the inverse: where sample k of local pixel lp lives
    __device__ __forceinline__ size_t slot_of(uint32_t lp, uint32_t k) const {
        if (sg_log2 == 0) return (size_t)k * n_local + lp;
        const uint32_t pl = 5u - sg_log2;
        const uint32_t batch = (k >> sg_log2) * n_blocks + (lp >> pl);
        return ((size_t)batch << 5) + ((k & ((1u << sg_log2) - 1u)) << pl) + (lp & ((1u << pl) - 1u));
    }
#endif
};

} // namespace paths
