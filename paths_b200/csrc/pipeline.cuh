// pipeline.cuh -- the wavefront kernels: raygen -> { extend -> shade -> shadow } x (max_bounces + 1) -> accumulate.
//
// Replaces the reference's thread pool (src/renderer.rs:34-110, src/worker.rs:43-87) and the per-path loop of
// trace_ray (src/trace.rs:7-121).  One pass advances n_slots camera paths in lock step; the records of live paths are
// 16-byte (f32) / 32-byte (f64) vectors kept DENSE BY QUEUE POSITION (dev_scene.h): the shade kernel compacts survivors
// and their rays into the next bounce's arrays (ballot + popc, staged per warp in shared memory, one atomicAdd per flush),
// so every kernel streams contiguous lines.  All kernels are persistent: a fixed grid of warps walks the queue, whose
// length is read from device memory, so no host round trip is needed inside a pass.
#pragma once
#include "../../include/paths_b200.h"
#include "dev_shade.cuh"
#include "dev_trace.cuh"

namespace paths {

#define PATHS_FULL_MASK 0xFFFFFFFFu
// resident 128-thread blocks per SM the register allocator must leave room for (65536 / (128 * regs))
#ifndef PATHS_LB_EXTEND
#define PATHS_LB_EXTEND 8
#endif
#ifndef PATHS_QBUF
/* per-warp staging of surviving paths (ray pair + throughput, 3 vectors each) in the shade kernel.  Sweep on B200 (profiles/r02f_variants.txt,
 * teapot / bunny M samples/s with two pass contexts): 96 -> 3718 / 2660, 64 -> 3773 / 2698, 48 -> 3400 / 2664: below 64 the queue-counter
 * atomics come back (one per ~24 survivors), above it the shade grid's shared memory (18 KB per block) shrinks L1 for the closest-hit
 * kernel of the other pass context that runs beside it */
#define PATHS_QBUF 64u
#endif
#ifndef PATHS_LB_SHADE
#define PATHS_LB_SHADE 5
#endif
#ifndef PATHS_LB_SHADE_PLAIN
#define PATHS_LB_SHADE_PLAIN 6 /* no lights, no extensions: the kernel without next-event estimation fits 6 blocks */
#endif
#ifndef PATHS_LB_SHADE64
#define PATHS_LB_SHADE64 4
#endif
#define PATHS_FLAG_SPECULAR 0x100u
#ifndef PATHS_SHADE_PARK
#define PATHS_SHADE_PARK 1
#endif
// The production instantiations of the traversal kernels (tuned on B200, profiles/): <rounds between refill votes, idle lanes that
// trigger a refill, inner nodes per round[, lanes that keep the inner phase going]>
#ifndef PATHS_E2_MIN
#define PATHS_E2_MIN 16
#endif
#ifndef PATHS_E2_KI
#define PATHS_E2_KI 8
#endif
#ifndef PATHS_E4_MIN
#define PATHS_E4_MIN 16
#endif
#ifndef PATHS_E4_KI
#define PATHS_E4_KI 16
#endif
#ifndef PATHS_E4_INNER
#define PATHS_E4_INNER 12
#endif
#define PATHS_K_EXTEND2 k_extend_refill<1, PATHS_E2_MIN, PATHS_E2_KI>
#define PATHS_K_SHADOW2 k_shadow_refill<1, 16, 8>
#define PATHS_K_EXTEND4 k_extend4s<1, PATHS_E4_MIN, PATHS_E4_KI, PATHS_E4_INNER>
#define PATHS_K_SHADOW4 k_shadow4<4, 16, 2>

template <typename R> struct IsF64 {
    static const bool value = false;
};
template <> struct IsF64<double> {
    static const bool value = true;
};

template <typename R> __device__ __forceinline__ typename Vec4T<R>::type ld4(const typename Vec4T<R>::type *p) { return *p; }
// Path records are read once and written once per kernel: streaming (evict-first) accesses keep them from pushing the scene
// tables out of L1/L2.
__device__ __forceinline__ float4 ld_rec(const float4 *p) { return __ldcs(p); }
__device__ __forceinline__ double4 ld_rec(const double4 *p) {
    double2 a = __ldcs(reinterpret_cast<const double2 *>(p)), b = __ldcs(reinterpret_cast<const double2 *>(p) + 1);
    return make_double4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void st_rec(float4 *p, float4 v) { __stcs(p, v); }
__device__ __forceinline__ void st_rec(double4 *p, double4 v) {
    __stcs(reinterpret_cast<double2 *>(p), make_double2(v.x, v.y));
    __stcs(reinterpret_cast<double2 *>(p) + 1, make_double2(v.z, v.w));
}

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ uint32_t warp_fetch(uint32_t *cursor) {
    uint32_t base = 0;
    if ((threadIdx.x & 31u) == 0) base = atomicAdd(cursor, 32u);
    return __shfl_sync(PATHS_FULL_MASK, base, 0);
}
// position in a queue for every lane with pred set; one atomic per warp
__device__ __forceinline__ uint32_t warp_append(uint32_t *counter, bool pred) {
    unsigned m = __ballot_sync(PATHS_FULL_MASK, pred);
    if (m == 0) return 0;
    unsigned lane = threadIdx.x & 31u;
    int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(PATHS_FULL_MASK, base, leader);
    return base + __popc(m & ((1u << lane) - 1u));
}

// ------------------------------------------------------------------ raygen (src/worker.rs:58-87, src/camera.rs:47-94)
template <typename R>
__global__ void __launch_bounds__(256) k_raygen(DevCamera cam, uint32_t k0, uint32_t k1, ShardInfo shard, uint32_t s_base, uint32_t n_slots,
                                                PathState<typename Vec4T<R>::type> ps, PassControl pc) {
    uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += stride) {
        uint32_t lp, k;
        shard.split(i, lp, k);
        uint32_t pixel = shard.pixel_of(lp);
        V3<R> o, d;
        R w;
        pixel_sample_ray<R>(cam, k0, k1, pixel, s_base + k, o, d, w);
        // bounce 0: queue position = slot; throughput (1, 1, 1) is implicit, "last bounce specular" starts true (trace.rs:8-11)
        st_rec(&ps.ray[0][2 * i], mk4(o.x, o.y, o.z, bits_to_real(i, R(0))));
        st_rec(&ps.ray[0][2 * i + 1], mk4(d.x, d.y, d.z, bits_to_real(PATHS_FLAG_SPECULAR, R(0))));
        st_rec(&ps.col[i], mk4(R(0), R(0), R(0), w));
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) pc.qcount[0] = n_slots;
}

// ------------------------------------------------------------------ extend: closest hit (src/scene.rs:172-197 -> src/bvh.rs:78-140)
template <typename R, bool MIXED>
__global__ void __launch_bounds__(128) k_extend(DevScene sc, const typename Vec4T<R>::type *__restrict__ rays, typename Vec4T<R>::type *__restrict__ hits,
                                                const uint32_t *count_ptr, uint32_t *cursor) {
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t base = warp_fetch(cursor);
        if (base >= count) break;
        uint32_t i = base + lane;
        if (i < count) {
            const uint32_t slot = i; // records are dense by queue position
            typename Vec4T<R>::type o = ld_rec(&rays[2 * i]), d = ld_rec(&rays[2 * i + 1]);
            if constexpr (IsF64<R>::value) {
                Hit64 h = MIXED ? trace_mixed<false>(sc, mk3((double)o.x, (double)o.y, (double)o.z), mk3((double)d.x, (double)d.y, (double)d.z),
                                                     (double)__int_as_float(0x7f800000), 0xFFFFFFFFu)
                                : trace_ref(sc, mk3((double)o.x, (double)o.y, (double)o.z), mk3((double)d.x, (double)d.y, (double)d.z));
                st_rec(&hits[slot], mk4(bits_to_real(h.prim, R(0)), R(h.t), R(h.bx), R(h.by)));
            } else {
                HitF h = trace_fast<false>(sc, (float)o.x, (float)o.y, (float)o.z, (float)d.x, (float)d.y, (float)d.z, __int_as_float(0x7f800000),
                                           0xFFFFFFFFu);
                st_rec(&hits[slot], mk4(bits_to_real(h.prim, R(0)), R(h.t), R(h.bx), R(h.by)));
            }
        }
    }
}

// ------------------------------------------------------------------ camera segment of the production path: raygen + closest hit
// Primary rays of neighbouring slots are nearly parallel, so a warp's 32 rays walk the tree together: plain batches of 32
// (no lane refill) are the faster schedule here (-2..3.5 % extend time), and generating the ray in the same kernel hides
// raygen's 64 B/slot of HBM writes behind the traversal and saves reading the ray back (32 B/slot).
// WIDE4 (scenes above 64k primitives): the closest hit walks the four-wide nodes (stack in shared memory) in a plain per-lane loop -- fewer
// dependent node fetches per ray, one 128-byte line per step.  Results are the same either way (closest hit, ties to the lower id).
template <typename R, bool WIDE4> // float only; a template so that both translation units may include this header
__global__ void __launch_bounds__(128, PATHS_LB_EXTEND) k_camera_extend(DevScene sc, DevCamera cam, uint32_t k0, uint32_t k1, ShardInfo shard,
                                                                        uint32_t s_base, uint32_t n_slots, PathState<float4> ps, PassControl pc,
                                                                        uint32_t *cursor) {
    extern __shared__ int32_t smem_stack[];
    const uint32_t lane = threadIdx.x & 31u;
    if (blockIdx.x == 0 && threadIdx.x == 0) pc.qcount[0] = n_slots;
    for (;;) {
        uint32_t base = warp_fetch(cursor);
        if (base >= n_slots) break;
        uint32_t i = base + lane;
        if (i < n_slots) {
            uint32_t lp, k;
            shard.split(i, lp, k);
            uint32_t pixel = shard.pixel_of(lp); // whole-frame passes: 32 consecutive local pixels are an 8 x 4 tile (ShardInfo)
            V3<float> o, d;
            float w;
            pixel_sample_ray<float>(cam, k0, k1, pixel, s_base + k, o, d, w);
            // bounce 0: queue position = slot; throughput (1, 1, 1) is implicit, "last bounce specular" starts true (trace.rs:8-11)
            st_rec(&ps.ray[0][2 * i], make_float4(o.x, o.y, o.z, __uint_as_float(i)));
            st_rec(&ps.ray[0][2 * i + 1], make_float4(d.x, d.y, d.z, __uint_as_float(PATHS_FLAG_SPECULAR)));
            st_rec(&ps.col[i], make_float4(0.0f, 0.0f, 0.0f, w));
            HitF h = WIDE4 ? trace_plain4(sc, o.x, o.y, o.z, d.x, d.y, d.z, smem_stack + threadIdx.x, blockDim.x)
                           : trace_fast<false>(sc, o.x, o.y, o.z, d.x, d.y, d.z, __int_as_float(0x7f800000), 0xFFFFFFFFu);
            st_rec(&ps.hit[i], make_float4(__uint_as_float(h.prim), h.t, h.bx, h.by));
        }
    }
}

// ------------------------------------------------------------------ which copy of the shading tables an instantiation reads
template <typename R> struct ShadeTables { // f64 (parity and mixed modes): the reference-precision tables
    typedef DevObject Object;
    static __device__ __forceinline__ void prim(const DevScene &sc, uint32_t prim, uint4 &meta, V3<R> &gvec) {
        meta = sc.prim_meta[prim];
        const double *g = sc.prim_geom + (size_t)prim * 12;
        gvec = (meta.x & PATHS_PM_TRI) ? mk3<R>(R(g[9]), R(g[10]), R(g[11])) : mk3<R>(R(g[0]), R(g[1]), R(g[2]));
    }
    static __device__ __forceinline__ const DevObject *objects(const DevScene &sc) { return sc.objects; }
    static __device__ __forceinline__ V3<R> vertex_normal(const DevScene &sc, size_t v) {
        const double *p = sc.vertex_normals + 3 * v;
        return mk3<R>(R(p[0]), R(p[1]), R(p[2]));
    }
    static __device__ __forceinline__ V3<R> vertex_colour(const DevScene &sc, size_t v) {
        const double *p = sc.vertex_colours + 3 * v;
        return mk3<R>(R(p[0]), R(p[1]), R(p[2]));
    }
};
template <> struct ShadeTables<float> { // production: f32 mirrors
    typedef DevObjectF Object;
    static __device__ __forceinline__ void prim(const DevScene &sc, uint32_t prim, uint4 &meta, V3<float> &gvec) {
        const float4 a = __ldg(sc.prim_shade + 2 * (size_t)prim), b = __ldg(sc.prim_shade + 2 * (size_t)prim + 1);
        meta = make_uint4(__float_as_uint(a.x), __float_as_uint(a.y), __float_as_uint(a.z), 0u);
        gvec = mk3<float>(b.x, b.y, b.z);
    }
    static __device__ __forceinline__ const DevObjectF *objects(const DevScene &sc) { return sc.objects32; }
    static __device__ __forceinline__ V3<float> vertex_normal(const DevScene &sc, size_t v) {
        const float4 n = __ldg(sc.vertex_normals32 + v);
        return mk3<float>(n.x, n.y, n.z);
    }
    static __device__ __forceinline__ V3<float> vertex_colour(const DevScene &sc, size_t v) {
        const float4 c = __ldg(sc.vertex_colours32 + v);
        return mk3<float>(c.x, c.y, c.z);
    }
};

// ------------------------------------------------------------------ shade: one iteration of trace_ray's loop body (src/trace.rs:13-118)
template <typename R> struct ShadeOut {
    bool alive, shadow, shadow2; // shadow2: environment NEE sample (extension)
    V3<R> sh_o, sh_d, sh_c, sh2_d, sh2_c;
    uint32_t light;
    // the surviving path's next segment (alive): origin, direction, throughput, loops | specular flag
    V3<R> next_o, next_d, next_thr;
    uint32_t next_flags;
};

// A ray that left the scene: the sky's radiance along it, weighted by the path's throughput, goes to the path's radiance record (trace.rs:20-23)
template <typename R, bool EXT>
__device__ __forceinline__ void shade_miss(const DevScene &sc, PathState<typename Vec4T<R>::type> &ps, V3<R> rd, V3<R> throughput, bool last_specular,
                                           uint32_t slot, uint32_t ext) {
    typedef typename Vec4T<R>::type R4;
    // with environment NEE the sky is accounted for at the previous vertex unless that bounce was specular (same rule
    // the reference applies to lights, trace.rs:31-41)
    if (EXT && (ext & PATHS_EXT_ENV_NEE) && sc.sky_type == 2 && !last_specular) return;
    R4 c4 = ps.col[slot]; // the radiance record is only touched by the branches that add light
    V3<R> colour = mk3<R>(c4.x, c4.y, c4.z);
    colour = colour + throughput * sky_light<R>(sc, neg(rd));
    ps.col[slot] = mk4(colour.x, colour.y, colour.z, c4.w);
}

// LIGHTS: the scene has lights (next-event estimation is compiled in); EXT: any estimator extension is on.  The common kernels --
// no extensions, with or without lights -- carry none of the other code and keep their registers for occupancy.
// o4 = {origin, slot}, d4 = {direction, flags}, h4 = hit record of this segment, throughput = the path's weight so far.
template <typename R, bool LIGHTS, bool EXT>
__device__ __forceinline__ void shade_path(const DevScene &sc, PathState<typename Vec4T<R>::type> &ps, typename Vec4T<R>::type o4,
                                           typename Vec4T<R>::type d4, typename Vec4T<R>::type h4, V3<R> throughput, uint32_t slot, uint32_t pixel,
                                           uint32_t S, uint32_t k0, uint32_t k1, uint32_t max_bounces, uint32_t ext, unsigned long long *totals,
                                           ShadeOut<R> &out) {
    typedef typename Vec4T<R>::type R4;
    V3<R> ro = mk3<R>(o4.x, o4.y, o4.z), rd = mk3<R>(d4.x, d4.y, d4.z);
    uint32_t flags = real_to_bits(d4.w);
    uint32_t loops = flags & 0xFFu;
    bool last_specular = (flags & PATHS_FLAG_SPECULAR) != 0;
    uint32_t prim = real_to_bits(h4.x);
    out.alive = false;
    out.shadow = false;
    out.shadow2 = false;
    const bool env_nee = EXT && (ext & PATHS_EXT_ENV_NEE) && sc.sky_type == 2;

    if (prim == 0xFFFFFFFFu) { // trace.rs:20-23
        shade_miss<R, EXT>(sc, ps, rd, throughput, last_specular, slot, ext);
        return;
    }
    // ---- Collision (geom.rs:59-70) + Scene::find_intersection post-processing (scene.rs:172-197)
    // The production instantiation reads f32 mirrors of the tables (the same values, rounded once on the host): one 32-byte record
    // per primitive instead of a 16-byte and a 96-byte one, float4 vertex normals, a float object record.
    typedef typename ShadeTables<R>::Object ObjectT;
    uint4 meta;
    V3<R> gvec; // triangle: rotated face normal; sphere: centre
    ShadeTables<R>::prim(sc, prim, meta, gvec);
    R t = h4.y, bx = h4.z, by = h4.w;
    V3<R> location = ro + (rd * t);
    V3<R> normal;
    bool is_tri = (meta.x & PATHS_PM_TRI) != 0;
    if (is_tri) { // geom.rs:293-302
        R cos_theta = dot(gvec, rd);
        normal = cos_theta > R(0) ? neg(gvec) : gvec;
    } else { // geom.rs:228-234
        normal = normed(location - gvec);
    }
    bool is_light = (meta.x & PATHS_PM_LIGHT) != 0;
    R bz = R(1.0) - bx - by;
    const ObjectT *obj = is_light ? nullptr : ShadeTables<R>::objects(sc) + meta.y;
    const uint32_t *face = nullptr;
    if (obj && obj->shape == 1 && is_tri) {
        face = sc.faces + 3 * ((size_t)obj->face_base + meta.z);
        if (obj->smooth) { // model.rs:142-156 (un-normalised, not flipped) then Mesh::rotate (geom.rs:119-121)
            V3<R> an = ShadeTables<R>::vertex_normal(sc, obj->vert_base + face[0]);
            V3<R> bn = ShadeTables<R>::vertex_normal(sc, obj->vert_base + face[1]);
            V3<R> cn = ShadeTables<R>::vertex_normal(sc, obj->vert_base + face[2]);
            V3<R> sn = an * bx + bn * by + cn * bz;
            normal = mat_apply<R>(obj->rot, sn);
        }
    }
    R cos_in = dot(rd, neg(normal)); // trace.rs:25-28
    if (cos_in <= R(0)) return;
    if (is_light) { // trace.rs:31-41
        if (last_specular) {
            const DevLight &l = sc.lights[meta.y];
            R4 c4 = ps.col[slot];
            V3<R> colour = mk3<R>(c4.x, c4.y, c4.z);
            colour = colour + throughput * mk3<R>(R(l.colour[0]), R(l.colour[1]), R(l.colour[2])) * R(l.intensity);
            if (col_bad(colour)) atomicAdd(totals + 2, 1ull);
            ps.col[slot] = mk4(colour.x, colour.y, colour.z, c4.w);
        }
        return;
    }
    // ---- material resolve (trace.rs:44-50, material.rs:39-49,183-195, model.rs:158-172)
    Material<R> mat = load_material<R>(*obj);
    mat.m.ggx = mat.diffuse.ggx = mat.specular.ggx = EXT && (ext & PATHS_EXT_GGX) != 0;
    const bool micro = EXT && (ext & (PATHS_EXT_MICROFACET | PATHS_EXT_MICROFACET_COSINE)) != 0 && (mat.fresnel || mat.m.type == 3);
    if (face && !mat.fresnel && (mat.m.type == 0 || mat.m.type == 2) && obj->vertex_albedo) {
        V3<R> ac = ShadeTables<R>::vertex_colour(sc, obj->vert_base + face[0]);
        V3<R> bc = ShadeTables<R>::vertex_colour(sc, obj->vert_base + face[1]);
        V3<R> cc = ShadeTables<R>::vertex_colour(sc, obj->vert_base + face[2]);
        mat.m.albedo = ac * bx + bc * by + cc * bz;
    }
    // block 0 = {light choice, light u, light v, gloss choice}; block 1 = {hemisphere u, v, roulette, -}: generated only when read
    U4 r0, r1;
    r0.v[0] = r0.v[1] = r0.v[2] = r0.v[3] = 0u;
    r1 = r0;
    if (LIGHTS || mat.m.type == 2 || micro) r0 = philox4x32_10(pixel, S, loops * 2u, PATHS_TAG_PATH, k0, k1);
    if (loops >= 2 || mat.m.type == 0 || mat.m.type == 2 || micro) r1 = philox4x32_10(pixel, S, loops * 2u + 1u, PATHS_TAG_PATH, k0, k1);
    V3<R> vec_out = neg(rd);
    V3<R> new_origin = location + normal * R(0.0001); // trace.rs:57,89

    // ---- next event estimation (trace.rs:53-82)
    if (LIGHTS) {
        uint32_t nl = sc.n_lights;
        uint32_t id = min(nl - 1u, (uint32_t)(u01<R>(r0.v[0]) * R(nl))); // scene.rs:199-206
        const DevLight &light = sc.lights[id];
        V3<R> in_dir;
        R inv_pdf;
        light_sample<R>(light, location, u01<R>(r0.v[1]), u01<R>(r0.v[2]), in_dir, inv_pdf);
        V3<R> sdir = neg(in_dir);
        R cos_theta = fmax_(R(0.0), dot(normal, sdir));
        if (cos_theta > R(0)) {
            V3<R> base = mk3<R>(R(light.colour[0]), R(light.colour[1]), R(light.colour[2])) * R(light.intensity);
            V3<R> brdf = material_brdf<R>(mat, vec_out, neg(sdir), normal);
            V3<R> direct = base * brdf * inv_pdf;
            if (EXT && (ext & PATHS_EXT_NEE_ALL_LIGHTS)) direct = direct * R(nl); // extension: undo the uniform light choice
            if (col_bad(direct)) atomicAdd(totals + 2, 1ull); // Colour::check (trace.rs:80)
            V3<R> contrib = direct * throughput;
            // production instantiation: a NaN contribution (the reference's Colour::check lets NaN through, colour.rs:56-60, and the
            // pixel stays NaN for the whole epoch) is counted as a bad sample and dropped; the f64 instantiation follows the reference
            bool usable = true;
            if constexpr (!IsF64<R>::value) {
                if (isnan_(contrib.x) || isnan_(contrib.y) || isnan_(contrib.z)) {
                    atomicAdd(totals + 2, 1ull);
                    usable = false;
                }
            }
            // a zero contribution cannot change the estimate: skip its visibility query
            if (usable && (contrib.x != R(0) || contrib.y != R(0) || contrib.z != R(0))) {
                out.shadow = true;
                out.sh_o = new_origin;
                out.sh_d = sdir;
                out.sh_c = contrib;
                out.light = id;
            }
        }
    }
    // ---- extension: one NEE sample toward the Hdri sky, drawn from its CDF table
    if (env_nee) {
        U4 r2 = philox4x32_10(pixel, S, loops, PATHS_TAG_ENV, k0, k1);
        V3<R> lookup, radiance;
        R epdf;
        env_sample<R>(sc, u01<R>(r2.v[0]), u01<R>(r2.v[1]), lookup, epdf, radiance);
        V3<R> sdir = neg(lookup); // a ray leaving along sdir sees ambient_light(-sdir) (trace.rs:21)
        if (dot(normal, sdir) > R(0) && epdf > R(0)) {
            V3<R> f = material_brdf<R>(mat, vec_out, lookup, normal); // vec_in = direction the light travels = -sdir
            V3<R> contrib = radiance * f * throughput / epdf;
            if (contrib.x != R(0) || contrib.y != R(0) || contrib.z != R(0)) {
                out.shadow2 = true;
                out.sh_o = new_origin;
                out.sh2_d = sdir;
                out.sh2_c = contrib;
            }
        }
    }
    // ---- BSDF sampling (trace.rs:84-98)
    V3<R> direction, brdf;
    R pdf;
    bool is_specular;
    if (micro) {
        if (!material_sample_ext<R>(mat, vec_out, normal, u01<R>(r0.v[0]), u01<R>(r0.v[3]), u01<R>(r1.v[0]), u01<R>(r1.v[1]),
                                    (ext & PATHS_EXT_MICROFACET_COSINE) != 0, direction, pdf, brdf, is_specular))
            return; // sampled below the horizon: zero weight
    } else if (!material_sample<R>(mat, vec_out, normal, u01<R>(r0.v[3]), u01<R>(r1.v[0]), u01<R>(r1.v[1]), direction, pdf, brdf, is_specular)) {
        atomicAdd(totals + 2, 1ull); // the reference panics: "Not implemented" (material.rs:86)
        return;
    }
    V3<R> attenuation = brdf / pdf;
    throughput = throughput * attenuation;
    if (cmax3(throughput) <= R(0)) return;
    // emittance is BLACK for every material the scene schema can build (serde.rs:278-291): trace.rs:100-101 adds zero
    if (loops >= 2) { // Russian roulette (trace.rs:104-111); survival chance is not clamped to 1
        R survival = cmax3(throughput);
        if (u01<R>(r1.v[2]) > survival) return;
        throughput = throughput / survival;
    }
    loops += 1;
    if (loops > max_bounces) return; // trace.rs:14-16
    out.next_o = new_origin;
    out.next_d = direction;
    out.next_thr = throughput;
    out.next_flags = loops | (is_specular ? PATHS_FLAG_SPECULAR : 0u);
    out.alive = true;
}

// Shade of bounce b: reads the dense records of parity b & 1 (ray pair, hit, throughput), writes the survivors' next segment densely
// into parity (b + 1) & 1 at the positions they take in the next queue.
template <typename R, bool LIGHTS, bool EXT>
__global__ void __launch_bounds__(128, (IsF64<R>::value ? PATHS_LB_SHADE64 : (LIGHTS || EXT) ? PATHS_LB_SHADE : PATHS_LB_SHADE_PLAIN)) k_shade(DevScene sc, PathState<typename Vec4T<R>::type> ps,
                                               const typename Vec4T<R>::type *__restrict__ rin, const typename Vec4T<R>::type *__restrict__ tin,
                                               typename Vec4T<R>::type *__restrict__ rout, typename Vec4T<R>::type *__restrict__ tout,
                                               ShardInfo shard, uint32_t s_base, uint32_t k0,
                                               uint32_t k1, uint32_t max_bounces, uint32_t ext, uint32_t bounce,
                                               const uint32_t *count_ptr, uint32_t *next_count, uint32_t *shadow_count, uint32_t *env_count,
                                               uint32_t env_base, uint32_t *cursor, unsigned long long *totals) {
    typedef typename Vec4T<R>::type R4;
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u;
    const bool first = bounce == 0; // queue position = slot, throughput (1, 1, 1) not stored
    // rin / tin: this bounce's dense records (parity bounce & 1); rout / tout: the next bounce's (the other parity).  Passed as kernel
    // parameters so that they are constant-bank operands, not registers (and never a dynamic index into the parameter block).
    // cursor == nullptr: batches of 32 entries are dealt to the warps of the grid round-robin (shading cost per entry is
    // uniform enough) -- no fetch atomic on the critical path; the next batch's records are prefetched into L2 meanwhile
    constexpr uint32_t QB = IsF64<R>::value ? 56u : PATHS_QBUF; // survivors staged per warp: {origin | slot, direction | flags} pairs, then throughputs
    __shared__ R4 qbuf_all[4][3 * QB];
    R4 *qbuf = qbuf_all[threadIdx.x >> 5];
    uint32_t fill = 0; // warp-uniform
    constexpr uint32_t SBUF = LIGHTS ? (IsF64<R>::value ? 56u : 96u) : 1u; // f64: 2 x 21.5 KB of static shared memory per block (48 KB limit) // shadow records staged per warp (origin | direction | contribution)
    __shared__ R4 sbuf_all[4][3 * SBUF];
    R4 *sbuf = sbuf_all[threadIdx.x >> 5];
    uint32_t sfill = 0; // warp-uniform
    const uint32_t warp_stride = gridDim.x * (blockDim.x >> 5) * 32u;
    uint32_t static_base = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 32u;
    // Parking pays where shading a hit is long (next-event estimation, extensions): bunny shade -12 %; the plain kernel (no lights, no
    // extensions) is too light for it: teapot shade +2 %, environment.yml +30 % (profiles/r02l_sweep.txt) -- it keeps plain batches.
    constexpr bool PARK = PATHS_SHADE_PARK && (LIGHTS || EXT);
    __shared__ uint32_t park_all[4][PARK ? 64 : 1];
    uint32_t *park = park_all[threadIdx.x >> 5];
    uint32_t npark = 0; // warp-uniform
    // one group of up to 32 queue entries: shading, then warp-level staging of shadow records and survivors
    auto shade_batch = [&](bool valid, uint32_t i) {
            ShadeOut<R> out;
            out.alive = out.shadow = out.shadow2 = false;
            uint32_t slot = 0;
            if (valid) {
                R4 o4 = rin[2 * (size_t)i], d4 = rin[2 * (size_t)i + 1], h4 = ps.hit[i];
                V3<R> thr = mk3<R>(R(1), R(1), R(1));
                if (!first) {
                    R4 t4 = tin[i];
                    thr = mk3<R>(t4.x, t4.y, t4.z);
                }
                slot = real_to_bits(o4.w);
                uint32_t lp, k;
                shard.split(slot, lp, k);
                uint32_t pixel = shard.pixel_of(lp);
                shade_path<R, LIGHTS, EXT>(sc, ps, o4, d4, h4, thr, slot, pixel, s_base + k, k0, k1, max_bounces, ext, totals, out);
            }
            if (LIGHTS) { // shadow records: staged per warp like the survivors below, appended SBUF - 32 or more at a time
                const unsigned m = __ballot_sync(PATHS_FULL_MASK, out.shadow);
                if (out.shadow) {
                    const uint32_t k = sfill + __popc(m & ((1u << lane) - 1u));
                    sbuf[k] = mk4(out.sh_o.x, out.sh_o.y, out.sh_o.z, bits_to_real(out.light, R(0)));
                    sbuf[SBUF + k] = mk4(out.sh_d.x, out.sh_d.y, out.sh_d.z, bits_to_real(slot, R(0)));
                    sbuf[2 * SBUF + k] = mk4(out.sh_c.x, out.sh_c.y, out.sh_c.z, R(0));
                }
                sfill += __popc(m);
                if (sfill > SBUF - 32u) {
                    __syncwarp();
                    uint32_t sbase = 0;
                    if (lane == 0) sbase = atomicAdd(shadow_count, sfill);
                    sbase = __shfl_sync(PATHS_FULL_MASK, sbase, 0);
                    PATHS_DASSERT(sbase + sfill <= env_base);
                    for (uint32_t k = lane; k < sfill; k += 32u) {
                        ps.sh_o[sbase + k] = sbuf[k];
                        ps.sh_d[sbase + k] = sbuf[SBUF + k];
                        ps.sh_c[sbase + k] = sbuf[2 * SBUF + k];
                    }
                    sfill = 0;
                    __syncwarp();
                }
            }
            if (EXT && (ext & PATHS_EXT_ENV_NEE)) { // warp-uniform
                // environment records go to the second half of the shadow arrays with a counter of their own; a launch of its own resolves
                // them after the light records (two records of one slot in one launch would race on the slot's radiance record)
                uint32_t epos = env_base + warp_append(env_count, out.shadow2);
                if (out.shadow2) {
                    ps.sh_o[epos] = mk4(out.sh_o.x, out.sh_o.y, out.sh_o.z, bits_to_real(PATHS_LIGHT_ENV, R(0)));
                    ps.sh_d[epos] = mk4(out.sh2_d.x, out.sh2_d.y, out.sh2_d.z, bits_to_real(slot, R(0)));
                    ps.sh_c[epos] = mk4(out.sh2_c.x, out.sh2_c.y, out.sh2_c.z, R(0));
                }
            }
            // surviving paths: their next segment is gathered per warp in shared memory and appended to the next bounce's dense arrays
            // QB - 32 or more at a time (one counter atomic and full-line stores per flush instead of one atomic per batch)
            {
                const unsigned m = __ballot_sync(PATHS_FULL_MASK, out.alive);
                if (out.alive) {
                    const uint32_t k = fill + __popc(m & ((1u << lane) - 1u));
                    qbuf[2 * k] = mk4(out.next_o.x, out.next_o.y, out.next_o.z, bits_to_real(slot, R(0)));
                    qbuf[2 * k + 1] = mk4(out.next_d.x, out.next_d.y, out.next_d.z, bits_to_real(out.next_flags, R(0)));
                    qbuf[2 * QB + k] = mk4(out.next_thr.x, out.next_thr.y, out.next_thr.z, R(0));
                }
                fill += __popc(m);
                if (fill > QB - 32u) {
                    __syncwarp();
                    uint32_t qbase = 0;
                    if (lane == 0) qbase = atomicAdd(next_count, fill);
                    qbase = __shfl_sync(PATHS_FULL_MASK, qbase, 0);
                    PATHS_DASSERT(qbase + fill <= env_base);
                    for (uint32_t k = lane; k < 2u * fill; k += 32u) rout[2 * (size_t)qbase + k] = qbuf[k];
                    for (uint32_t k = lane; k < fill; k += 32u) tout[qbase + k] = qbuf[2 * QB + k];
                    fill = 0;
                    __syncwarp();
                }
            }
    };
    for (;;) {
        uint32_t base;
        if (cursor) base = warp_fetch(cursor);
        else {
            base = static_base;
            static_base += warp_stride;
            uint32_t in = base + warp_stride + lane;
            if (in < count && base < count) {
                prefetch_l2(rin + 2 * (size_t)in); // {origin, direction}: one 32-byte sector (f32)
                if (IsF64<R>::value) prefetch_l2(rin + 2 * (size_t)in + 1);
                prefetch_l2(ps.hit + in);
                if (!first) prefetch_l2(tin + in);
            }
        }
        uint32_t i = base + lane;
        const bool in_range = i < count;
        if constexpr (PARK) {
        // f64: ONE call site for shade_batch -- the last, partial group of parked hits goes through the loop too.  (With a second call
        // after the loop the f64 instantiation was not inlined: its closure and the kernel's parameter block went to a 928-byte local
        // frame per thread, which does not fit L1 beside the staging buffers -- profiles/r02zz_doom_shade_*.)  The f32 instantiations
        // inline both calls and keep the tail after the loop (one register less in the loop).
        bool valid = true;
        if (base >= count) {
            if (!IsF64<R>::value || npark == 0) break;
            __syncwarp();
            valid = lane < npark;
            i = valid ? park[lane] : 0u;
            npark = 0; // the next fetch is past the end again, and nothing is parked: the loop ends
        } else {
        // Rays that left the scene are finished at once; hits are PARKED (their queue positions, per warp, in shared memory) and shaded 32
        // at a time, so that the long material / sampling code runs with full warps (ncu on teapot bounce 1, where 45 % of the rays
        // miss: 15 of 32 lanes per instruction before).  Records are dense and prefetched, so the parked gather costs no extra latency.
        bool is_hit = false;
        if (in_range) {
            const R4 h4 = ps.hit[i];
            if (real_to_bits(h4.x) == 0xFFFFFFFFu) {
                const R4 d4 = rin[2 * (size_t)i + 1];
                V3<R> thr = mk3<R>(R(1), R(1), R(1));
                if (!first) {
                    const R4 t4 = tin[i];
                    thr = mk3<R>(t4.x, t4.y, t4.z);
                }
                const uint32_t mslot = first ? i : real_to_bits(rin[2 * (size_t)i].w);
                shade_miss<R, EXT>(sc, ps, mk3<R>(d4.x, d4.y, d4.z), thr, (real_to_bits(d4.w) & PATHS_FLAG_SPECULAR) != 0, mslot, ext);
            } else is_hit = true;
        }
        {
            const unsigned pm = __ballot_sync(PATHS_FULL_MASK, is_hit);
            if (is_hit) park[npark + __popc(pm & ((1u << lane) - 1u))] = i;
            npark += __popc(pm);
        }
        if (npark < 32u) continue;
        __syncwarp();
        i = park[lane];
        const uint32_t left = npark - 32u; // < 32
        __syncwarp();
        const uint32_t moved = lane < left ? park[32u + lane] : 0u;
        __syncwarp();
        if (lane < left) park[lane] = moved;
        npark = left;
        }
        shade_batch(valid, i);
        } else {
            if (base >= count) break;
            shade_batch(in_range, i);
        }
    }
    if constexpr (PARK && !IsF64<R>::value) {
        if (npark > 0) { // the last, partial group of parked hits
            __syncwarp();
            shade_batch(lane < npark, lane < npark ? park[lane] : 0u);
        }
    }
    if (LIGHTS && sfill > 0) {
        __syncwarp();
        uint32_t sbase = 0;
        if (lane == 0) sbase = atomicAdd(shadow_count, sfill);
        sbase = __shfl_sync(PATHS_FULL_MASK, sbase, 0);
        PATHS_DASSERT(sbase + sfill <= env_base);
        for (uint32_t k = lane; k < sfill; k += 32u) {
            ps.sh_o[sbase + k] = sbuf[k];
            ps.sh_d[sbase + k] = sbuf[SBUF + k];
            ps.sh_c[sbase + k] = sbuf[2 * SBUF + k];
        }
    }
    if (fill > 0) {
        __syncwarp();
        uint32_t qbase = 0;
        if (lane == 0) qbase = atomicAdd(next_count, fill);
        qbase = __shfl_sync(PATHS_FULL_MASK, qbase, 0);
        PATHS_DASSERT(qbase + fill <= env_base);
        for (uint32_t k = lane; k < 2u * fill; k += 32u) rout[2 * (size_t)qbase + k] = qbuf[k];
        for (uint32_t k = lane; k < fill; k += 32u) tout[qbase + k] = qbuf[2 * QB + k];
    }
}

// ------------------------------------------------------------------ shadow: visibility of the NEE sample (src/trace.rs:61-66)
template <typename R, bool MIXED>
__global__ void __launch_bounds__(128) k_shadow(DevScene sc, PathState<typename Vec4T<R>::type> ps, const uint32_t *count_ptr, uint32_t *cursor,
                                                unsigned long long *totals) {
    typedef typename Vec4T<R>::type R4;
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t base = warp_fetch(cursor);
        if (base >= count) break;
        uint32_t i = base + lane;
        if (i < count) {
            R4 o = ld_rec(&ps.sh_o[i]), d = ld_rec(&ps.sh_d[i]);
            uint32_t lid = real_to_bits(o.w), slot = real_to_bits(d.w);
            bool occluded;
            if constexpr (IsF64<R>::value && MIXED) {
                // any-hit over the production tree with f64 primitives: the same predicate as below, no rounding to f32
                double t_light = (double)__int_as_float(0x7f800000);
                uint32_t skip = 0xFFFFFFFFu;
                D3v so = mk3((double)o.x, (double)o.y, (double)o.z), sdir = mk3((double)d.x, (double)d.y, (double)d.z);
                if (lid != PATHS_LIGHT_ENV) {
                    const DevLight &l = sc.lights[lid];
                    skip = l.prim;
                    double tl;
                    if (l.geometry == 1 && sphere_hit64(mk3(l.v[0], l.v[1], l.v[2]), l.radius, so, sdir, tl)) t_light = tl;
                }
                occluded = trace_mixed<true>(sc, so, sdir, t_light, skip).prim != 0xFFFFFFFFu;
            } else if constexpr (IsF64<R>::value) {
                // the reference runs a full closest-hit query and compares entities (trace.rs:61-66)
                Hit64 h = trace_ref(sc, mk3((double)o.x, (double)o.y, (double)o.z), mk3((double)d.x, (double)d.y, (double)d.z));
                occluded = false;
                if (h.prim != 0xFFFFFFFFu) {
                    uint4 m = sc.prim_meta[h.prim];
                    occluded = lid == PATHS_LIGHT_ENV || !((m.x & PATHS_PM_LIGHT) && m.y == lid);
                }
            } else {
                // same predicate as an any-hit query: occluded <=> some primitive of another entity lies closer than the
                // nearest hit on the light's own sphere (t_light = inf for point lights, which own no primitive)
                float t_light = __int_as_float(0x7f800000);
                uint32_t skip = 0xFFFFFFFFu;
                if (lid != PATHS_LIGHT_ENV) {
                    const DevLight &l = sc.lights[lid];
                    skip = l.prim;
                    if (l.geometry == 1) {
                        double tl;
                        if (sphere_hit64(mk3(l.v[0], l.v[1], l.v[2]), l.radius, mk3((double)o.x, (double)o.y, (double)o.z),
                                         mk3((double)d.x, (double)d.y, (double)d.z), tl))
                            t_light = (float)tl;
                    }
                }
                HitF h = trace_fast<true>(sc, (float)o.x, (float)o.y, (float)o.z, (float)d.x, (float)d.y, (float)d.z, t_light, skip);
                occluded = h.prim != 0xFFFFFFFFu;
            }
            if (!occluded) { // trace.rs:81: colour += direct * throughput
                R4 c = ld_rec(&ps.sh_c[i]);
                R4 col = ld_rec(&ps.col[slot]);
                col.x += c.x;
                col.y += c.y;
                col.z += c.z;
                if (!(col.x >= R(0) && col.y >= R(0) && col.z >= R(0))) atomicAdd(totals + 2, 1ull); // Colour::check (trace.rs:82); NaN counts
                st_rec(&ps.col[slot], col);
            }
        }
    }
}

// ------------------------------------------------------------------ production extend / shadow with lane refill (f32 only)
// Persistent warps; a lane whose ray has ended takes the next queue entry as soon as MIN_REFILL lanes of its warp are idle,
// instead of waiting for the slowest ray of a batch of 32.  ROUNDS bounded traversal rounds run between two refill checks;
// one round = up to three inner nodes, then at most one leaf.
template <int ROUNDS, int MIN_REFILL, int KI = 3>
__global__ void __launch_bounds__(128, PATHS_LB_EXTEND) k_extend_refill(DevScene sc, const float4 *__restrict__ rays, float4 *__restrict__ hits,
                                                       const uint32_t *count_ptr, uint32_t *cursor) {
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes = sc.fast_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    int32_t stack[PATHS_FAST_STACK];
    TravF tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t slot = 0;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    slot = i; // the hit record goes to the ray's own queue position
                    float4 o = ld_rec(&rays[2 * (size_t)i]), d = ld_rec(&rays[2 * (size_t)i + 1]);
                    tv.init(o.x, o.y, o.z, d.x, d.y, d.z, __int_as_float(0x7f800000), sc.fast_root);
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
#pragma unroll 1
        for (int r = 0; r < ROUNDS; r++) {
            if (has) {
#pragma unroll 1
                for (int k = 0; k < KI && tv.node >= 0; k++) tv.inner(nodes, stack);
                if (tv.node < 0 && tv.node != PATHS_TRAV_DONE) tv.leaf<false>(leaves, stack, 0xFFFFFFFFu);
                if (tv.node == PATHS_TRAV_DONE) {
                    st_rec(&hits[slot], make_float4(__uint_as_float(tv.prim), tv.t, tv.bx, tv.by));
                    has = false;
                }
            }
        }
    }
}

template <int ROUNDS, int MIN_REFILL, int KI = 3>
__global__ void __launch_bounds__(128, PATHS_LB_EXTEND) k_shadow_refill(DevScene sc, PathState<float4> ps, const uint32_t *count_ptr, uint32_t *cursor,
                                                       unsigned long long *totals) {
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes = sc.fast_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    int32_t stack[PATHS_FAST_STACK];
    TravF tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t rec = 0, slot = 0, skip = 0xFFFFFFFFu;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    rec = i;
                    float4 o = ld_rec(&ps.sh_o[i]), d = ld_rec(&ps.sh_d[i]);
                    uint32_t lid = __float_as_uint(o.w);
                    slot = __float_as_uint(d.w);
                    float t_light = __int_as_float(0x7f800000);
                    skip = 0xFFFFFFFFu;
                    if (lid != PATHS_LIGHT_ENV) { // nearest hit on the light's own sphere bounds the search (see k_shadow)
                        const DevLight &l = sc.lights[lid];
                        skip = l.prim;
                        if (l.geometry == 1) {
                            double tl;
                            if (sphere_hit64(mk3(l.v[0], l.v[1], l.v[2]), l.radius, mk3((double)o.x, (double)o.y, (double)o.z),
                                             mk3((double)d.x, (double)d.y, (double)d.z), tl))
                                t_light = (float)tl;
                        }
                    }
                    tv.init(o.x, o.y, o.z, d.x, d.y, d.z, t_light, sc.fast_root);
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
#pragma unroll 1
        for (int r = 0; r < ROUNDS; r++) {
            if (has) {
#pragma unroll 1
                for (int k = 0; k < KI && tv.node >= 0; k++) tv.inner(nodes, stack);
                bool occluded = false;
                if (tv.node < 0 && tv.node != PATHS_TRAV_DONE) occluded = tv.leaf<true>(leaves, stack, skip);
                if (occluded) has = false;
                else if (tv.node == PATHS_TRAV_DONE) {
                    float4 c = ld_rec(&ps.sh_c[rec]);
                    float4 col = ld_rec(&ps.col[slot]);
                    col.x += c.x; col.y += c.y; col.z += c.z;
                    if (!(col.x >= 0.0f && col.y >= 0.0f && col.z >= 0.0f)) atomicAdd(totals + 2, 1ull);
                    st_rec(&ps.col[slot], col);
                    has = false;
                }
            }
        }
    }
}

// ------------------------------------------------------------------ mixed mode with lane refill (f64 records, TravM)
// The plain batches of k_extend<double, true> leave a warp waiting for its slowest ray (doom.yml at 1080p: 95 of 189 ms per 32 spp in
// extend).  Same schedule as k_extend_refill over the f64 path state; every ray's answer is bit-identical to trace_mixed's.
#ifndef PATHS_M2_MIN
#define PATHS_M2_MIN 8 /* refill as soon as 8 lanes idle: an f64 leaf is long, idle lanes cost more than in f32 (doom +1.8 % over 16; profiles/r02zz_mixed_refill_ab.txt) */
#endif
#ifndef PATHS_M2_KI
#define PATHS_M2_KI 8
#endif
#ifndef PATHS_LB_M2
#define PATHS_LB_M2 6
#endif
#ifndef PATHS_M2_SPEC
#define PATHS_M2_SPEC 0 /* > 0: k_extend2s_m with this MIN_INNER (parked leaves) */
#endif
#if PATHS_M2_SPEC > 0
#define PATHS_K_EXTEND2M k_extend2s_m<PATHS_M2_MIN, PATHS_M2_KI, PATHS_M2_SPEC>
#else
#define PATHS_K_EXTEND2M k_extend_refill_m<PATHS_M2_MIN, PATHS_M2_KI>
#endif
#define PATHS_K_SHADOW2M k_shadow_refill_m<PATHS_M2_MIN, PATHS_M2_KI>
template <int MIN_REFILL, int KI>
__global__ void __launch_bounds__(128, PATHS_LB_M2) k_extend_refill_m(DevScene sc, const double4 *__restrict__ rays, double4 *__restrict__ hits,
                                                         const uint32_t *count_ptr, uint32_t *cursor) {
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes = sc.fast_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    int32_t stack[PATHS_FAST_STACK];
    TravM tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t slot = 0;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    slot = i;
                    double4 o = ld_rec(&rays[2 * (size_t)i]), d = ld_rec(&rays[2 * (size_t)i + 1]);
                    tv.init<false>(mk3(o.x, o.y, o.z), mk3(d.x, d.y, d.z), (double)__int_as_float(0x7f800000), sc.fast_root);
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
        if (has) {
#pragma unroll 1
            for (int k = 0; k < KI && tv.node >= 0; k++) tv.inner(nodes, stack);
            if (tv.node < 0 && tv.node != PATHS_TRAV_DONE) tv.leaf<false>(sc, leaves, stack, 0xFFFFFFFFu);
            if (tv.node == PATHS_TRAV_DONE) {
                st_rec(&hits[slot], mk4(bits_to_real(tv.prim, 0.0), tv.t, tv.bx, tv.by));
                has = false;
            }
        }
    }
}

// Mixed mode, parked leaves: a lane that reaches a leaf parks it and keeps descending from its stack; the leaves are tested when fewer
// than MIN_INNER lanes still have an inner node to visit, so that the long f64 primitive tests run with many lanes at once (ncu on
// doom.yml bounce 1 with k_extend_refill_m: the triangle test ran at 7 of 32 lanes, the slab tests at 15).  Same answers (TravM::leaf_test).
template <int MIN_REFILL, int KI, int MIN_INNER>
__global__ void __launch_bounds__(128) k_extend2s_m(DevScene sc, const double4 *__restrict__ rays, double4 *__restrict__ hits,
                                                    const uint32_t *count_ptr, uint32_t *cursor) {
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes = sc.fast_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    int32_t stack[PATHS_FAST_STACK];
    TravM tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t slot = 0;
    int32_t pend = 0;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    slot = i;
                    double4 o = ld_rec(&rays[2 * (size_t)i]), d = ld_rec(&rays[2 * (size_t)i + 1]);
                    tv.init<false>(mk3(o.x, o.y, o.z), mk3(d.x, d.y, d.z), (double)__int_as_float(0x7f800000), sc.fast_root);
                    pend = 0;
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        const int n_has = __popc(__ballot_sync(PATHS_FULL_MASK, has));
        if (n_has == 0) break;
#pragma unroll 1
        for (int k = 0; k < KI; k++) {
            bool inner = has && tv.node >= 0;
            int n_inner = __popc(__ballot_sync(PATHS_FULL_MASK, inner));
            if (n_inner == 0 || (n_inner < MIN_INNER && n_inner < n_has)) break;
            if (inner) tv.inner(nodes, stack);
            if (has && pend == 0 && tv.node < 0 && tv.node != PATHS_TRAV_DONE) {
                pend = tv.node;
                tv.node = tv.pop(stack);
            }
        }
        if (has) {
            if (pend == 0 && tv.node < 0 && tv.node != PATHS_TRAV_DONE) { // reached while the inner phase was cut short
                pend = tv.node;
                tv.node = tv.pop(stack);
            }
            if (pend != 0) {
                tv.leaf_test<false>(sc, leaves, pend, 0xFFFFFFFFu);
                pend = 0;
            }
            if (tv.node == PATHS_TRAV_DONE) {
                st_rec(&hits[slot], mk4(bits_to_real(tv.prim, 0.0), tv.t, tv.bx, tv.by));
                has = false;
            }
        }
    }
}

template <int MIN_REFILL, int KI>
__global__ void __launch_bounds__(128, PATHS_LB_M2) k_shadow_refill_m(DevScene sc, PathState<double4> ps, const uint32_t *count_ptr, uint32_t *cursor,
                                                         unsigned long long *totals) {
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes = sc.fast_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    int32_t stack[PATHS_FAST_STACK];
    TravM tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t rec = 0, slot = 0, skip = 0xFFFFFFFFu;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    rec = i;
                    double4 o = ld_rec(&ps.sh_o[i]), d = ld_rec(&ps.sh_d[i]);
                    uint32_t lid = real_to_bits(o.w);
                    slot = real_to_bits(d.w);
                    double t_light = (double)__int_as_float(0x7f800000);
                    skip = 0xFFFFFFFFu;
                    D3v so = mk3(o.x, o.y, o.z), sdir = mk3(d.x, d.y, d.z);
                    if (lid != PATHS_LIGHT_ENV) { // nearest hit on the light's own sphere bounds the search (see k_shadow)
                        const DevLight &l = sc.lights[lid];
                        skip = l.prim;
                        double tl;
                        if (l.geometry == 1 && sphere_hit64(mk3(l.v[0], l.v[1], l.v[2]), l.radius, so, sdir, tl)) t_light = tl;
                    }
                    tv.init<true>(so, sdir, t_light, sc.fast_root);
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
        if (has) {
#pragma unroll 1
            for (int k = 0; k < KI && tv.node >= 0; k++) tv.inner(nodes, stack);
            bool occluded = false;
            if (tv.node < 0 && tv.node != PATHS_TRAV_DONE) occluded = tv.leaf<true>(sc, leaves, stack, skip);
            if (occluded) has = false;
            else if (tv.node == PATHS_TRAV_DONE) { // trace.rs:81: colour += direct * throughput
                double4 c = ld_rec(&ps.sh_c[rec]);
                double4 col = ld_rec(&ps.col[slot]);
                col.x += c.x; col.y += c.y; col.z += c.z;
                if (!(col.x >= 0.0 && col.y >= 0.0 && col.z >= 0.0)) atomicAdd(totals + 2, 1ull); // Colour::check (trace.rs:82); NaN counts
                st_rec(&ps.col[slot], col);
                has = false;
            }
        }
    }
}

// Mixed mode over the four-wide nodes (scenes above 64k primitives, like the f32 path): lane refill, stack of 3 * depth + 1 levels per
// thread in dynamic shared memory.  Leaves are tested where they are met (parking them lost in the two-wide mixed kernel).
#ifndef PATHS_M4_KI
#define PATHS_M4_KI 4
#endif
#define PATHS_K_EXTEND4M k_extend4_m<PATHS_M2_MIN, PATHS_M4_KI>
#define PATHS_K_SHADOW4M k_shadow4_m<PATHS_M2_MIN, PATHS_M4_KI>
template <int MIN_REFILL, int KI>
__global__ void __launch_bounds__(128, PATHS_LB_M2) k_extend4_m(DevScene sc, const double4 *__restrict__ rays, double4 *__restrict__ hits,
                                                                  const uint32_t *count_ptr, uint32_t *cursor) {
    extern __shared__ int32_t smem_stack[];
    int32_t *sstack = smem_stack + threadIdx.x;
    const uint32_t stride = blockDim.x;
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes4 = sc.fast4_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    TravM4 tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t slot = 0;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    slot = i;
                    double4 o = ld_rec(&rays[2 * (size_t)i]), d = ld_rec(&rays[2 * (size_t)i + 1]);
                    tv.init<false>(mk3(o.x, o.y, o.z), mk3(d.x, d.y, d.z), (double)__int_as_float(0x7f800000), sc.fast4_root);
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
        if (has) {
#pragma unroll 1
            for (int k = 0; k < KI && tv.node >= 0; k++) tv.inner4<true>(nodes4, sstack, stride);
            if (tv.node < 0 && tv.node != PATHS_TRAV_DONE) {
                tv.leaf_test<false>(sc, leaves, tv.node, 0xFFFFFFFFu);
                tv.node = tv.pop4(sstack, stride);
            }
            if (tv.node == PATHS_TRAV_DONE) {
                st_rec(&hits[slot], mk4(bits_to_real(tv.prim, 0.0), tv.t, tv.bx, tv.by));
                has = false;
            }
        }
    }
}

template <int MIN_REFILL, int KI>
__global__ void __launch_bounds__(128, PATHS_LB_M2) k_shadow4_m(DevScene sc, PathState<double4> ps, const uint32_t *count_ptr, uint32_t *cursor,
                                                                  unsigned long long *totals) {
    extern __shared__ int32_t smem_stack[];
    int32_t *sstack = smem_stack + threadIdx.x;
    const uint32_t stride = blockDim.x;
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes4 = sc.fast4_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    TravM4 tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t rec = 0, slot = 0, skip = 0xFFFFFFFFu;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    rec = i;
                    double4 o = ld_rec(&ps.sh_o[i]), d = ld_rec(&ps.sh_d[i]);
                    uint32_t lid = real_to_bits(o.w);
                    slot = real_to_bits(d.w);
                    double t_light = (double)__int_as_float(0x7f800000);
                    skip = 0xFFFFFFFFu;
                    D3v so = mk3(o.x, o.y, o.z), sdir = mk3(d.x, d.y, d.z);
                    if (lid != PATHS_LIGHT_ENV) { // nearest hit on the light's own sphere bounds the search (see k_shadow)
                        const DevLight &l = sc.lights[lid];
                        skip = l.prim;
                        double tl;
                        if (l.geometry == 1 && sphere_hit64(mk3(l.v[0], l.v[1], l.v[2]), l.radius, so, sdir, tl)) t_light = tl;
                    }
                    tv.init<true>(so, sdir, t_light, sc.fast4_root);
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
        if (has) {
#pragma unroll 1
            for (int k = 0; k < KI && tv.node >= 0; k++) tv.inner4<false>(nodes4, sstack, stride);
            bool occluded = false;
            if (tv.node < 0 && tv.node != PATHS_TRAV_DONE) {
                occluded = tv.leaf_test<true>(sc, leaves, tv.node, skip);
                if (!occluded) tv.node = tv.pop4(sstack, stride);
            }
            if (occluded) has = false;
            else if (tv.node == PATHS_TRAV_DONE) { // trace.rs:81: colour += direct * throughput
                double4 c = ld_rec(&ps.sh_c[rec]);
                double4 col = ld_rec(&ps.col[slot]);
                col.x += c.x; col.y += c.y; col.z += c.z;
                if (!(col.x >= 0.0 && col.y >= 0.0 && col.z >= 0.0)) atomicAdd(totals + 2, 1ull); // Colour::check (trace.rs:82); NaN counts
                st_rec(&ps.col[slot], col);
                has = false;
            }
        }
    }
}

// ------------------------------------------------------------------ four-wide extend / shadow (lane refill, shared-memory stack)
template <int ROUNDS, int MIN_REFILL, int KI>
__global__ void __launch_bounds__(128, PATHS_LB_EXTEND) k_shadow4(DevScene sc, PathState<float4> ps, const uint32_t *count_ptr, uint32_t *cursor,
                                                                  unsigned long long *totals) {
    extern __shared__ int32_t smem_stack[];
    int32_t *sstack = smem_stack + threadIdx.x;
    const uint32_t stride = blockDim.x;
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes4 = sc.fast4_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    TravF4 tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t rec = 0, slot = 0, skip = 0xFFFFFFFFu;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    rec = i;
                    float4 o = ld_rec(&ps.sh_o[i]), d = ld_rec(&ps.sh_d[i]);
                    uint32_t lid = __float_as_uint(o.w);
                    slot = __float_as_uint(d.w);
                    float t_light = __int_as_float(0x7f800000);
                    skip = 0xFFFFFFFFu;
                    if (lid != PATHS_LIGHT_ENV) { // nearest hit on the light's own sphere bounds the search (see k_shadow)
                        const DevLight &l = sc.lights[lid];
                        skip = l.prim;
                        if (l.geometry == 1) {
                            double tl;
                            if (sphere_hit64(mk3(l.v[0], l.v[1], l.v[2]), l.radius, mk3((double)o.x, (double)o.y, (double)o.z),
                                             mk3((double)d.x, (double)d.y, (double)d.z), tl))
                                t_light = (float)tl;
                        }
                    }
                    tv.init(o.x, o.y, o.z, d.x, d.y, d.z, t_light, sc.fast4_root);
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
#pragma unroll 1
        for (int r = 0; r < ROUNDS; r++) {
            if (has) {
#pragma unroll 1
                for (int k = 0; k < KI && tv.node >= 0; k++) tv.inner4<false>(nodes4, sstack, stride);
                bool occluded = false;
                if (tv.node < 0 && tv.node != PATHS_TRAV_DONE) {
                    occluded = tv.leaf_test<true>(leaves, tv.node, skip);
                    tv.node = tv.pop4(sstack, stride);
                }
                if (occluded) has = false;
                else if (tv.node == PATHS_TRAV_DONE) {
                    float4 c = ld_rec(&ps.sh_c[rec]);
                    float4 col = ld_rec(&ps.col[slot]);
                    col.x += c.x; col.y += c.y; col.z += c.z;
                    if (!(col.x >= 0.0f && col.y >= 0.0f && col.z >= 0.0f)) atomicAdd(totals + 2, 1ull);
                    st_rec(&ps.col[slot], col);
                    has = false;
                }
            }
        }
    }
}

// Four-wide, speculative: a lane that reaches a leaf parks it (`pend`) and keeps descending from its stack; the inner phase of
// a round ends when fewer than MIN_INNER lanes still have an inner node to visit (one ballot per step), so that the leaf phase
// runs with most lanes of the warp (ncu on k_extend4: triangle tests ran with 3.6 of 32 lanes, box tests with 13.8).
template <int ROUNDS, int MIN_REFILL, int KI, int MIN_INNER>
__global__ void __launch_bounds__(128, PATHS_LB_EXTEND) k_extend4s(DevScene sc, const float4 *__restrict__ rays, float4 *__restrict__ hits,
                                                                   const uint32_t *count_ptr, uint32_t *cursor) {
    extern __shared__ int32_t smem_stack[];
    int32_t *sstack = smem_stack + threadIdx.x;
    const uint32_t stride = blockDim.x;
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes4 = sc.fast4_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    TravF4 tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t slot = 0;
    int32_t pend = 0;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    slot = i; // the hit record goes to the ray's own queue position
                    float4 o = ld_rec(&rays[2 * (size_t)i]), d = ld_rec(&rays[2 * (size_t)i + 1]);
                    tv.init(o.x, o.y, o.z, d.x, d.y, d.z, __int_as_float(0x7f800000), sc.fast4_root);
                    pend = 0;
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
#pragma unroll 1
        for (int r = 0; r < ROUNDS; r++) {
            const int n_has = __popc(__ballot_sync(PATHS_FULL_MASK, has));
#pragma unroll 1
            for (int k = 0; k < KI; k++) {
                bool inner = has && tv.node >= 0;
                // leave the inner phase when nobody is at an inner node, or when few are and other lanes wait with a leaf
                int n_inner = __popc(__ballot_sync(PATHS_FULL_MASK, inner));
                if (n_inner == 0 || (n_inner < MIN_INNER && n_inner < n_has)) break;
                if (inner) tv.inner4(nodes4, sstack, stride);
                if (has && pend == 0 && tv.node < 0 && tv.node != PATHS_TRAV_DONE) {
                    pend = tv.node;
                    tv.node = tv.pop4(sstack, stride);
                }
            }
            if (has) {
                if (pend == 0 && tv.node < 0 && tv.node != PATHS_TRAV_DONE) { // reached while the inner phase was cut short
                    pend = tv.node;
                    tv.node = tv.pop4(sstack, stride);
                }
                if (pend != 0) {
                    tv.leaf_test<false>(leaves, pend, 0xFFFFFFFFu);
                    pend = 0;
                }
                if (tv.node == PATHS_TRAV_DONE) {
                    st_rec(&hits[slot], make_float4(__uint_as_float(tv.prim), tv.t, tv.bx, tv.by));
                    has = false;
                }
            }
        }
    }
}

// Two-wide, speculative: the schedule of k_extend4s over two-wide nodes and a local stack.  A lane that reaches a leaf parks it and keeps
// descending from its stack; the inner phase ends when fewer than MIN_INNER lanes still have an inner node to visit.
template <int ROUNDS, int MIN_REFILL, int KI, int MIN_INNER>
__global__ void __launch_bounds__(128, PATHS_LB_EXTEND) k_extend2s(DevScene sc, const float4 *__restrict__ rays, float4 *__restrict__ hits,
                                                                   const uint32_t *count_ptr, uint32_t *cursor) {
    const uint32_t count = *count_ptr;
    const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes = sc.fast_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    int32_t stack[PATHS_FAST_STACK];
    TravF tv;
    tv.node = PATHS_TRAV_DONE;
    bool has = false, exhausted = false;
    uint32_t slot = 0;
    int32_t pend = 0;
    for (;;) {
        unsigned idle = __ballot_sync(PATHS_FULL_MASK, !has);
        int n = __popc(idle);
        if (!exhausted && n >= MIN_REFILL) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(cursor, (uint32_t)n);
            base = __shfl_sync(PATHS_FULL_MASK, base, leader);
            if (!has) {
                uint32_t i = base + __popc(idle & lt);
                if (i < count) {
                    slot = i;
                    float4 o = ld_rec(&rays[2 * (size_t)i]), d = ld_rec(&rays[2 * (size_t)i + 1]);
                    tv.init(o.x, o.y, o.z, d.x, d.y, d.z, __int_as_float(0x7f800000), sc.fast_root);
                    pend = 0;
                    has = true;
                }
            }
            exhausted = base + (uint32_t)n >= count;
        }
        if (!__any_sync(PATHS_FULL_MASK, has)) break;
#pragma unroll 1
        for (int r = 0; r < ROUNDS; r++) {
            const int n_has = __popc(__ballot_sync(PATHS_FULL_MASK, has));
#pragma unroll 1
            for (int k = 0; k < KI; k++) {
                bool inner = has && tv.node >= 0;
                int n_inner = __popc(__ballot_sync(PATHS_FULL_MASK, inner));
                if (n_inner == 0 || (n_inner < MIN_INNER && n_inner < n_has)) break;
                if (inner) tv.inner(nodes, stack);
                if (has && pend == 0 && tv.node < 0 && tv.node != PATHS_TRAV_DONE) {
                    pend = tv.node;
                    tv.node = tv.pop(stack);
                }
            }
            if (has) {
                if (pend == 0 && tv.node < 0 && tv.node != PATHS_TRAV_DONE) { // reached while the inner phase was cut short
                    pend = tv.node;
                    tv.node = tv.pop(stack);
                }
                if (pend != 0) {
                    tv.leaf_test<false>(leaves, pend, 0xFFFFFFFFu);
                    pend = 0;
                }
                if (tv.node == PATHS_TRAV_DONE) {
                    st_rec(&hits[slot], make_float4(__uint_as_float(tv.prim), tv.t, tv.bx, tv.by));
                    has = false;
                }
            }
        }
    }
}

// ------------------------------------------------------------------ accumulate (src/pixels.rs:6-31,49-51; worker.rs:76 applies the camera weight)
template <typename R>
__global__ void __launch_bounds__(256) k_accumulate(PathState<typename Vec4T<R>::type> ps, ShardInfo shard, uint32_t spp_pass,
                                                    typename Vec4T<R>::type *sums) {
    typedef typename Vec4T<R>::type R4;
    uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t lp = blockIdx.x * blockDim.x + threadIdx.x; lp < shard.n_local; lp += stride) {
        uint32_t pixel = shard.pixel_of(lp);
        R4 s = sums[pixel];
        for (uint32_t k = 0; k < spp_pass; k++) {
            R4 c = ld_rec(&ps.col[shard.slot_of(lp, k)]);
            s.x += c.x * c.w;
            s.y += c.y * c.w;
            s.z += c.z * c.w;
        }
        s.w += R(spp_pass);
        sums[pixel] = s;
    }
}

template <typename R> __global__ void k_tally(PassControl pc, uint32_t n_bounces, uint32_t n_slots) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        unsigned long long e = 0, s = 0;
        for (uint32_t b = 0; b < n_bounces; b++) {
            e += pc.qcount[b];
            s += pc.scount[b] + pc.ecount[b];
        }
        atomicAdd(pc.totals + 0, e); // two pass contexts may tally at the same time
        atomicAdd(pc.totals + 1, s);
        atomicAdd(pc.totals + 3, (unsigned long long)n_slots);
    }
}

} // namespace paths
