// dev_trace.cuh -- ray / primitive intersection and BVH traversal on the device.
//
//   * trace_ref<>   f64, the reference's binary tree and visiting order (src/bvh.rs:78-140): bit-exact ids and distances.
//   * trace_fast<>  production: f32 slab tests over 64-byte two-child nodes, f32 Moeller-Trumbore triangles,
//                   f64 spheres (the scenes' "planes" are spheres of radius 1e6: (l.oc)^2 - oc.oc + r^2 cancels ~1e12
//                   against ~1e12, hopeless in f32), optional any-hit early-out for shadow rays.
#pragma once
#include "dev_math.cuh"
#include "dev_scene.h"

namespace paths {

typedef V3<double> D3v;

struct Hit64 {
    uint32_t prim;
    double t, bx, by;
};

// ------------------------------------------------------------------ primitives, f64, reference operation order
// src/geom.rs:208-235 (assumes a unit direction)
__device__ __forceinline__ bool sphere_hit64(D3v c, double r, D3v o, D3v l, double &t) {
    D3v oc = o - c;
    double loc = dot(l, oc);
    double disc = (loc * loc) - dot(oc, oc) + (r * r);
    if (disc < 0.0) return false;
    double tmp = -loc;
    double sq = sqrt(disc);
    double d1 = tmp + sq;
    double d2 = tmp - sq;
    if (d1 < 0.0) return false;
    t = d2 > 0.0 ? d2 : d1;
    return true;
}

// src/geom.rs:265-303
__device__ __forceinline__ bool triangle_hit64(const double *g, D3v o, D3v dir, double &t, double &bx, double &by) {
    D3v a = mk3(g[0], g[1], g[2]), b = mk3(g[3], g[4], g[5]), c = mk3(g[6], g[7], g[8]), n = mk3(g[9], g[10], g[11]);
    double cos_theta = dot(n, dir);
    double d = dot(n, a);
    double tt = (d - dot(n, o)) / cos_theta;
    if (isnan(tt) || tt < 0.0) return false;
    D3v p = o + (dir * tt);
    double area_abc = dot(n, cross(b - a, c - a));
    double area_pbc = dot(n, cross(b - p, c - p));
    double area_pca = dot(n, cross(c - p, a - p));
    double x = area_pbc / area_abc;
    double y = area_pca / area_abc;
    double z = 1.0 - x - y;
    if (x < 0.0 || y < 0.0 || z < 0.0) return false;
    t = tt;
    bx = x;
    by = y;
    return true;
}

__device__ __forceinline__ bool prim_hit64(const DevScene &sc, uint32_t prim, D3v o, D3v d, double &t, double &bx, double &by) {
    const double *g = sc.prim_geom + (size_t)prim * 12;
    if (sc.prim_meta[prim].x & PATHS_PM_TRI) return triangle_hit64(g, o, d, t, bx, by);
    bx = 0.0;
    by = 0.0;
    return sphere_hit64(mk3(g[0], g[1], g[2]), g[3], o, d, t);
}

// src/bvh.rs:8-21: strict tmin < tmax, optional tmin < len, no tmax >= 0 test; Rust f64::min/max ignore NaN like fmin/fmax
__device__ __forceinline__ bool slab_ref(const double *box, D3v o, D3v inv, bool has_len, double len, double &tmin_out) {
    D3v t0 = (mk3(box[0], box[1], box[2]) - o) * inv;
    D3v t1 = (mk3(box[3], box[4], box[5]) - o) * inv;
    D3v tsm = mk3(fmin(t0.x, t1.x), fmin(t0.y, t1.y), fmin(t0.z, t1.z));
    D3v tbg = mk3(fmax(t0.x, t1.x), fmax(t0.y, t1.y), fmax(t0.z, t1.z));
    double tmin = vmax3(tsm);
    double tmax = vmin3(tbg);
    if (tmin < tmax && (!has_len || tmin < len)) {
        tmin_out = tmin;
        return true;
    }
    return false;
}

#define PATHS_REF_STACK 100 /* src/bvh.rs:79 */

// src/bvh.rs:78-140.  Returns the closest hit in the reference's visiting order (first found wins exact ties).
static __device__ __noinline__ Hit64 trace_ref(const DevScene &sc, D3v o, D3v d) {
    Hit64 best;
    best.prim = 0xFFFFFFFFu;
    best.t = 0.0;
    best.bx = best.by = 0.0;
    D3v inv = mk3(1.0 / d.x, 1.0 / d.y, 1.0 / d.z); // geom.rs:39-46, vector.rs:63-65
    double tmp;
    if (!slab_ref(sc.ref_root_box, o, inv, false, 0.0, tmp)) return best;
    int32_t stack[PATHS_REF_STACK];
    int sp = 0;
    int32_t node = sc.ref_root;
    for (;;) {
        if (node < 0) { // leaf (:92-110)
            uint32_t prim = (uint32_t)~node;
            double t, bx, by;
            if (prim_hit64(sc, prim, o, d, t, bx, by)) {
                if (best.prim == 0xFFFFFFFFu || t < best.t) {
                    best.prim = prim;
                    best.t = t;
                    best.bx = bx;
                    best.by = by;
                }
            }
            if (sp == 0) break;
            node = stack[--sp];
        } else { // cluster (:111-137)
            const DevRefNode &nd = sc.ref_nodes[node];
            bool has = best.prim != 0xFFFFFFFFu;
            double ld = 0.0, rd = 0.0;
            bool lh = slab_ref(nd.lbox, o, inv, has, best.t, ld);
            bool rh = slab_ref(nd.rbox, o, inv, has, best.t, rd);
            if (lh && rh) {
                if (sp >= PATHS_REF_STACK) break; // the reference would panic (index out of bounds)
                if (ld < rd) {
                    stack[sp++] = nd.right;
                    node = nd.left;
                } else {
                    stack[sp++] = nd.left;
                    node = nd.right;
                }
            } else if (lh) node = nd.left;
            else if (rh) node = nd.right;
            else {
                if (sp == 0) break;
                node = stack[--sp];
            }
        }
    }
    return best;
}

// ------------------------------------------------------------------ production traversal
struct HitF {
    uint32_t prim;
    float t, bx, by;
};

#define PATHS_FAST_STACK 64

// ANY = true: return as soon as any primitive other than `skip_prim` is hit closer than tmax (shadow rays).
template <bool ANY> __device__ __forceinline__ HitF trace_fast(const DevScene &sc, float ox, float oy, float oz, float dx, float dy, float dz,
                                                                float tmax, uint32_t skip_prim) {
    HitF best;
    best.prim = 0xFFFFFFFFu;
    best.t = tmax;
    best.bx = best.by = 0.0f;
    const float eps = 8.27180613e-25f; // 2^-80: keeps 1/d finite for axis-parallel rays
    float idx = 1.0f / (fabsf(dx) > eps ? dx : copysignf(eps, dx));
    float idy = 1.0f / (fabsf(dy) > eps ? dy : copysignf(eps, dy));
    float idz = 1.0f / (fabsf(dz) > eps ? dz : copysignf(eps, dz));
    float oix = ox * idx, oiy = oy * idy, oiz = oz * idz;
    int32_t stack[PATHS_FAST_STACK];
    int sp = 0;
    int32_t node = sc.fast_root;
    const float4 *__restrict__ nodes = sc.fast_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    for (;;) {
        while (node >= 0) {
            const float4 n0 = __ldg(nodes + 4 * node);     // c0: lo.x hi.x lo.y hi.y
            const float4 n1 = __ldg(nodes + 4 * node + 1); // c1: lo.x hi.x lo.y hi.y
            const float4 n2 = __ldg(nodes + 4 * node + 2); // c0 lo.z hi.z, c1 lo.z hi.z
            const float4 n3 = __ldg(nodes + 4 * node + 3); // child codes
            float a0 = n0.x * idx - oix, a1 = n0.y * idx - oix, a2 = n0.z * idy - oiy, a3 = n0.w * idy - oiy;
            float a4 = n2.x * idz - oiz, a5 = n2.y * idz - oiz;
            float tn0 = fmaxf(fmaxf(fminf(a0, a1), fminf(a2, a3)), fmaxf(fminf(a4, a5), 0.0f));
            float tf0 = fminf(fminf(fmaxf(a0, a1), fmaxf(a2, a3)), fminf(fmaxf(a4, a5), best.t)) * 1.0000004f;
            float b0 = n1.x * idx - oix, b1 = n1.y * idx - oix, b2 = n1.z * idy - oiy, b3 = n1.w * idy - oiy;
            float b4 = n2.z * idz - oiz, b5 = n2.w * idz - oiz;
            float tn1 = fmaxf(fmaxf(fminf(b0, b1), fminf(b2, b3)), fmaxf(fminf(b4, b5), 0.0f));
            float tf1 = fminf(fminf(fmaxf(b0, b1), fmaxf(b2, b3)), fminf(fmaxf(b4, b5), best.t)) * 1.0000004f;
            bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
            int32_t c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (h0 && h1) {
                bool first0 = tn0 <= tn1;
                node = first0 ? c0 : c1;
                if (sp < PATHS_FAST_STACK) stack[sp++] = first0 ? c1 : c0;
            } else if (h0) node = c0;
            else if (h1) node = c1;
            else {
                if (sp == 0) return best;
                node = stack[--sp];
            }
        }
        // leaf: up to 8 slots of 3 x float4
        {
            uint32_t code = (uint32_t)~node;
            uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
            for (uint32_t k = 0; k < cnt; k++) {
                const float4 s0 = __ldg(leaves + 3 * (first + k));
                const float4 s1 = __ldg(leaves + 3 * (first + k) + 1);
                const float4 s2 = __ldg(leaves + 3 * (first + k) + 2);
                uint32_t tag = __float_as_uint(s2.w);
                uint32_t prim = tag & 0x7FFFFFFFu;
                if (ANY && prim == skip_prim) continue;
                if (tag & 0x80000000u) { // sphere: centre.xy | centre.z radius | -, -, -, tag   (f64 pairs)
                    double cx = __hiloint2double(__float_as_int(s0.y), __float_as_int(s0.x));
                    double cy = __hiloint2double(__float_as_int(s0.w), __float_as_int(s0.z));
                    double cz = __hiloint2double(__float_as_int(s1.y), __float_as_int(s1.x));
                    double rr = __hiloint2double(__float_as_int(s1.w), __float_as_int(s1.z));
                    double t;
                    if (sphere_hit64(mk3(cx, cy, cz), rr, mk3((double)ox, (double)oy, (double)oz), mk3((double)dx, (double)dy, (double)dz), t)) {
                        float tf = (float)t;
                        if (tf < best.t) {
                            best.t = tf;
                            best.prim = prim;
                            best.bx = best.by = 0.0f;
                            if (ANY) return best;
                        }
                    }
                } else { // triangle: v0 | e1 | e2, tag   (Moeller-Trumbore)
                    float e1x = s1.x, e1y = s1.y, e1z = s1.z, e2x = s2.x, e2y = s2.y, e2z = s2.z;
                    float px = dy * e2z - dz * e2y, py = dz * e2x - dx * e2z, pz = dx * e2y - dy * e2x;
                    float det = e1x * px + e1y * py + e1z * pz;
                    float inv = 1.0f / det;
                    float tx = ox - s0.x, ty = oy - s0.y, tz = oz - s0.z;
                    float u = (tx * px + ty * py + tz * pz) * inv;
                    float qx = ty * e1z - tz * e1y, qy = tz * e1x - tx * e1z, qz = tx * e1y - ty * e1x;
                    float v = (dx * qx + dy * qy + dz * qz) * inv;
                    float t = (e2x * qx + e2y * qy + e2z * qz) * inv;
                    if (u >= 0.0f && v >= 0.0f && u + v <= 1.0f && t >= 0.0f && t < best.t) {
                        best.t = t;
                        best.prim = prim;
                        best.bx = 1.0f - u - v; // weight of vertex a (geom.rs:289-291)
                        best.by = u;            // weight of vertex b
                        if (ANY) return best;
                    }
                }
            }
            if (sp == 0) return best;
            node = stack[--sp];
        }
    }
}

// ------------------------------------------------------------------ resumable production traversal (persistent warps with lane refill)
// The same arithmetic as trace_fast, cut into bounded steps so that a warp can hand a finished lane a new ray instead of
// idling until its slowest ray ends (ncu on the batch-of-32 version: 7-9 of 32 threads active per instruction on bounce rays).
#define PATHS_TRAV_DONE ((int32_t)0x80000000)

struct TravF {
    float ox, oy, oz, dx, dy, dz, idx, idy, idz, oix, oiy, oiz;
    float t, bx, by;
    uint32_t prim;
    int32_t node;
    int sp;

    __device__ __forceinline__ void init(float ox_, float oy_, float oz_, float dx_, float dy_, float dz_, float tmax, int32_t root) {
        ox = ox_; oy = oy_; oz = oz_; dx = dx_; dy = dy_; dz = dz_;
        const float eps = 8.27180613e-25f;
        idx = 1.0f / (fabsf(dx) > eps ? dx : copysignf(eps, dx));
        idy = 1.0f / (fabsf(dy) > eps ? dy : copysignf(eps, dy));
        idz = 1.0f / (fabsf(dz) > eps ? dz : copysignf(eps, dz));
        oix = ox * idx; oiy = oy * idy; oiz = oz * idz;
        t = tmax; bx = by = 0.0f;
        prim = 0xFFFFFFFFu;
        node = root;
        sp = 0;
    }
    // one inner node: picks the nearer hit child, pushes the other, or pops
    __device__ __forceinline__ void inner(const float4 *__restrict__ nodes, int32_t *stack) {
        const float4 n0 = __ldg(nodes + 4 * node);
        const float4 n1 = __ldg(nodes + 4 * node + 1);
        const float4 n2 = __ldg(nodes + 4 * node + 2);
        const float4 n3 = __ldg(nodes + 4 * node + 3);
        float a0 = n0.x * idx - oix, a1 = n0.y * idx - oix, a2 = n0.z * idy - oiy, a3 = n0.w * idy - oiy;
        float a4 = n2.x * idz - oiz, a5 = n2.y * idz - oiz;
        float tn0 = fmaxf(fmaxf(fminf(a0, a1), fminf(a2, a3)), fmaxf(fminf(a4, a5), 0.0f));
        float tf0 = fminf(fminf(fmaxf(a0, a1), fmaxf(a2, a3)), fminf(fmaxf(a4, a5), t)) * 1.0000004f;
        float b0 = n1.x * idx - oix, b1 = n1.y * idx - oix, b2 = n1.z * idy - oiy, b3 = n1.w * idy - oiy;
        float b4 = n2.z * idz - oiz, b5 = n2.w * idz - oiz;
        float tn1 = fmaxf(fmaxf(fminf(b0, b1), fminf(b2, b3)), fmaxf(fminf(b4, b5), 0.0f));
        float tf1 = fminf(fminf(fmaxf(b0, b1), fmaxf(b2, b3)), fminf(fmaxf(b4, b5), t)) * 1.0000004f;
        bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
        int32_t c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
        if (h0 && h1) {
            bool first0 = tn0 <= tn1;
            node = first0 ? c0 : c1;
            if (sp < PATHS_FAST_STACK) stack[sp++] = first0 ? c1 : c0;
        } else if (h0) node = c0;
        else if (h1) node = c1;
        else node = sp > 0 ? stack[--sp] : PATHS_TRAV_DONE;
    }
    // one leaf (node < 0, not DONE): tests its primitives, then pops.  Returns true when ANY and something was hit.
    template <bool ANY> __device__ __forceinline__ bool leaf(const float4 *__restrict__ leaves, int32_t *stack, uint32_t skip_prim) {
        uint32_t code = (uint32_t)~node;
        uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
        for (uint32_t k = 0; k < cnt; k++) {
            const float4 s0 = __ldg(leaves + 3 * (first + k));
            const float4 s1 = __ldg(leaves + 3 * (first + k) + 1);
            const float4 s2 = __ldg(leaves + 3 * (first + k) + 2);
            uint32_t tag = __float_as_uint(s2.w);
            uint32_t p = tag & 0x7FFFFFFFu;
            if (ANY && p == skip_prim) continue;
            if (tag & 0x80000000u) {
                double cx = __hiloint2double(__float_as_int(s0.y), __float_as_int(s0.x));
                double cy = __hiloint2double(__float_as_int(s0.w), __float_as_int(s0.z));
                double cz = __hiloint2double(__float_as_int(s1.y), __float_as_int(s1.x));
                double rr = __hiloint2double(__float_as_int(s1.w), __float_as_int(s1.z));
                double tt;
                if (sphere_hit64(mk3(cx, cy, cz), rr, mk3((double)ox, (double)oy, (double)oz), mk3((double)dx, (double)dy, (double)dz), tt)) {
                    float tf = (float)tt;
                    if (tf < t) {
                        t = tf; prim = p; bx = by = 0.0f;
                        if (ANY) return true;
                    }
                }
            } else {
                float e1x = s1.x, e1y = s1.y, e1z = s1.z, e2x = s2.x, e2y = s2.y, e2z = s2.z;
                float px = dy * e2z - dz * e2y, py = dz * e2x - dx * e2z, pz = dx * e2y - dy * e2x;
                float det = e1x * px + e1y * py + e1z * pz;
                float inv = 1.0f / det;
                float tx = ox - s0.x, ty = oy - s0.y, tz = oz - s0.z;
                float u = (tx * px + ty * py + tz * pz) * inv;
                float qx = ty * e1z - tz * e1y, qy = tz * e1x - tx * e1z, qz = tx * e1y - ty * e1x;
                float v = (dx * qx + dy * qy + dz * qz) * inv;
                float tt = (e2x * qx + e2y * qy + e2z * qz) * inv;
                if (u >= 0.0f && v >= 0.0f && u + v <= 1.0f && tt >= 0.0f && tt < t) {
                    t = tt; prim = p; bx = 1.0f - u - v; by = u;
                    if (ANY) return true;
                }
            }
        }
        node = sp > 0 ? stack[--sp] : PATHS_TRAV_DONE;
        return false;
    }
};

} // namespace paths
