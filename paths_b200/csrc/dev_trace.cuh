// dev_trace.cuh -- ray / primitive intersection and BVH traversal on the device.
//
//   * trace_ref<>   f64, the reference's binary tree and visiting order (src/bvh.rs:78-140): bit-exact ids and distances.
//   * trace_fast<>  production: f32 slab tests over 64-byte two-child nodes, watertight f32 triangles,
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

// the same test for a primitive met in a production-tree leaf: the leaf tag's top bit says "sphere" (bvh_fast.cpp), which saves the
// dependent load of prim_meta on the traversal path
__device__ __forceinline__ bool prim_hit64_tagged(const DevScene &sc, uint32_t tag, D3v o, D3v d, double &t, double &bx, double &by) {
    const double *g = sc.prim_geom + (size_t)(tag & 0x7FFFFFFFu) * 12;
    if (!(tag & 0x80000000u)) return triangle_hit64(g, o, d, t, bx, by);
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


// ------------------------------------------------------------------ watertight f32 ray / triangle test
// Woop, Benthin, Wald 2013, with the axis permutation folded into two shear vectors p, q (both perpendicular to d):
// x' = p . (v - o), y' = q . (v - o) put every vertex into the ray's own 2-D frame, where coordinates are of the size of
// the triangle, not of its distance; edge functions U = Cx' By' - Cy' Bx' etc. are then evaluated with separately rounded
// products (__fmul_rn, never contracted).  A vertex shared by two triangles gets bit-identical x', y' in both, so the
// shared edge's function comes out EXACTLY negated: a ray cannot slip between two triangles that share an edge.
// (Moeller-Trumbore in f32 lets ~3e-4 of the rays through an 871k-triangle mesh; the naive d . (B' x C') form is
// watertight too but its barycentrics are off by eps * dist^2 / edge^2 -- 4e-2 on that mesh.)
// Barycentrics are the reference's area ratios: bx = area(p,b,c) / area(a,b,c) (src/geom.rs:284-291).
struct RayFrame {
    float px, py, pz, qx, qy, qz;
};
__device__ __forceinline__ RayFrame ray_frame(float dx, float dy, float dz) {
    // Branch-free: lanes of a warp disagree on the dominant axis after the first bounce, and three divergent arms (a division each) cost
    // three times the warp instructions of one arm plus selects.  Same arithmetic per lane as the three-armed form: -u / dk, -v / dk.
    RayFrame f;
    const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    const bool kz_z = az >= ax && az >= ay, kz_y = !kz_z && ay >= ax;
    const float ik = 1.0f / (kz_z ? dz : (kz_y ? dy : dx));
    const float pu = -(kz_z ? dx : (kz_y ? dz : dy)) * ik; // kz = z: p = (1, 0, -dx/dz)   kz = y: p = (0, -dz/dy, 1)   kz = x: p = (-dy/dx, 1, 0)
    const float qv = -(kz_z ? dy : (kz_y ? dx : dz)) * ik; //         q = (0, 1, -dy/dz)           q = (1, -dx/dy, 0)           q = (-dz/dx, 0, 1)
    f.px = kz_z ? 1.0f : (kz_y ? 0.0f : pu);
    f.py = kz_z ? 0.0f : (kz_y ? pu : 1.0f);
    f.pz = kz_z ? pu : (kz_y ? 1.0f : 0.0f);
    f.qx = kz_z ? 0.0f : (kz_y ? 1.0f : qv);
    f.qy = kz_z ? 1.0f : (kz_y ? qv : 0.0f);
    f.qz = kz_z ? qv : (kz_y ? 0.0f : 1.0f);
    return f;
}
struct TriHit {
    float t, bx, by;
};
__device__ __forceinline__ bool triangle_hit_wt(float4 s0, float4 s1, float4 s2, const RayFrame &f, float ox, float oy, float oz, float dx, float dy,
                                                float dz, float tmax, TriHit &h) {
    float a0 = s0.x - ox, a1 = s0.y - oy, a2 = s0.z - oz;
    float b0 = s1.x - ox, b1 = s1.y - oy, b2 = s1.z - oz;
    float c0 = s2.x - ox, c1 = s2.y - oy, c2 = s2.z - oz;
    // one fixed evaluation order for every vertex (explicit fma chain): a shared vertex must map to the same x', y' in both triangles
    float Ax = __fmaf_rn(f.pz, a2, __fmaf_rn(f.py, a1, __fmul_rn(f.px, a0))), Ay = __fmaf_rn(f.qz, a2, __fmaf_rn(f.qy, a1, __fmul_rn(f.qx, a0)));
    float Bx = __fmaf_rn(f.pz, b2, __fmaf_rn(f.py, b1, __fmul_rn(f.px, b0))), By = __fmaf_rn(f.qz, b2, __fmaf_rn(f.qy, b1, __fmul_rn(f.qx, b0)));
    float Cx = __fmaf_rn(f.pz, c2, __fmaf_rn(f.py, c1, __fmul_rn(f.px, c0))), Cy = __fmaf_rn(f.qz, c2, __fmaf_rn(f.qy, c1, __fmul_rn(f.qx, c0)));
    float U = __fmul_rn(Cx, By) - __fmul_rn(Cy, Bx);
    float V = __fmul_rn(Ax, Cy) - __fmul_rn(Ay, Cx);
    float W = __fmul_rn(Bx, Ay) - __fmul_rn(By, Ax);
    float mn = fminf(U, fminf(V, W)), mx = fmaxf(U, fmaxf(V, W));
    if (mn < 0.0f && mx > 0.0f) return false;
    float det = U + V + W;
    if (det == 0.0f) return false;
    // distance from the plane equation with n = (b - a) x (c - a): its error grows like eps / cos(theta) only
    float e1x = s1.x - s0.x, e1y = s1.y - s0.y, e1z = s1.z - s0.z, e2x = s2.x - s0.x, e2y = s2.y - s0.y, e2z = s2.z - s0.z;
    float nx = e1y * e2z - e1z * e2y, ny = e1z * e2x - e1x * e2z, nz = e1x * e2y - e1y * e2x;
    float t = (nx * a0 + ny * a1 + nz * a2) / (nx * dx + ny * dy + nz * dz);
    if (!(t >= 0.0f && t <= tmax)) return false; // t == tmax passes: the caller breaks exact ties by primitive id
    float inv = 1.0f / det;
    h.t = t;
    h.bx = U * inv;
    h.by = V * inv;
    return true;
}

// ------------------------------------------------------------------ production traversal
// Closest hit with an ORDER-INDEPENDENT rule for exact ties: of two primitives hit at bit-equal distances the lower id wins.  The box
// tests never cull a candidate at t == best (tf = min(far, best) * 1.0000004 >= best), so every production kernel -- plain batches,
// two-wide lane refill, four-wide with parked leaves -- returns the same (prim, t) for the same ray whatever order it visits the tree
// in (tests/test_gpu_production_rays.py asserts that bit for bit).  The reference takes the first hit in ITS visiting order
// (src/bvh.rs:93-104), which the f64 instantiation reproduces; a visibility query (any hit) is unaffected.
__device__ __forceinline__ bool closer(float t_new, uint32_t prim_new, float t_best, uint32_t prim_best) {
    return t_new < t_best || (t_new == t_best && prim_new < prim_best);
}
struct HitF {
    uint32_t prim;
    float t, bx, by;
};

#define PATHS_FAST_STACK 64
#ifndef PATHS_INNER_SELECT
#define PATHS_INNER_SELECT 0 /* two-wide inner node: child choice with selects and a predicated push instead of a four-way branch */
#endif
// A two-wide traversal holds at most one postponed sibling per tree level, and both builders refuse trees deeper than 60 levels
// (bvh_fast.cpp: error; bvh_device.cu: falls back to the host builder), so the 64-entry stack cannot overflow.  The release kernels keep a
// predicated store (an atomic in this cold branch cost 8-12 % of the closest-hit kernels' time: measured, profiles/r02c_ab.txt); the
// debug-assert build (make EXTRA=-DPATHS_DEBUG_ASSERTS) counts every dropped push and every failed device assertion, and the ray-level
// tests run once under it (paths_stats.stack_overflows / assert_failures, live only in that build).  One copy per translation unit.
static __device__ unsigned long long g_stack_overflows = 0ull;
static __device__ unsigned long long g_assert_failures = 0ull;
#ifdef PATHS_DEBUG_ASSERTS
#define PATHS_DASSERT(cond)                                   \
    do {                                                      \
        if (!(cond)) atomicAdd(&g_assert_failures, 1ull);     \
    } while (0)
#else
#define PATHS_DASSERT(cond) ((void)0)
#endif
__device__ __forceinline__ void stack_push(int32_t *stack, int &sp, int32_t v) {
    if (sp < PATHS_FAST_STACK) stack[sp++] = v;
#ifdef PATHS_DEBUG_ASSERTS
    else atomicAdd(&g_stack_overflows, 1ull);
#endif
}

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
    const RayFrame frame = ray_frame(dx, dy, dz);
    int32_t stack[PATHS_FAST_STACK];
    int sp = 0;
    int32_t node = sc.fast_root;
    if (node == (int32_t)0x80000000) return best; // empty tree
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
#if PATHS_INNER_SELECT
            const bool both = h0 && h1, any = h0 || h1, first0 = tn0 <= tn1;
            const int32_t near = (h0 && (!h1 || first0)) ? c0 : c1, far = first0 ? c1 : c0;
            if (both) stack_push(stack, sp, far);
            node = near;
            if (!any) {
                if (sp == 0) return best;
                node = stack[--sp];
            }
#else
            if (h0 && h1) {
                bool first0 = tn0 <= tn1;
                node = first0 ? c0 : c1;
                stack_push(stack, sp, first0 ? c1 : c0);
            } else if (h0) node = c0;
            else if (h1) node = c1;
            else {
                if (sp == 0) return best;
                node = stack[--sp];
            }
#endif
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
                        if (ANY ? tf < best.t : closer(tf, prim, best.t, best.prim)) {
                            best.t = tf;
                            best.prim = prim;
                            best.bx = best.by = 0.0f;
                            if (ANY) return best;
                        }
                    }
                } else { // triangle: v0 | v1 | v2, tag
                    TriHit th;
                    if (triangle_hit_wt(s0, s1, s2, frame, ox, oy, oz, dx, dy, dz, best.t, th) && (ANY ? th.t < best.t : closer(th.t, prim, best.t, best.prim))) {
                        best.t = th.t;
                        best.prim = prim;
                        best.bx = th.bx;
                        best.by = th.by;
                        if (ANY) return best;
                    }
                }
            }
            if (sp == 0) return best;
            node = stack[--sp];
        }
    }
}

// ------------------------------------------------------------------ mixed traversal: production tree, f64 primitives
// For scenes whose coordinates are too large for f32 rays (the integrator offsets new origins by an absolute 1e-4, trace.rs:57,89)
// but which should not pay for the reference's own tree: f32 slab tests over the production SAH tree with the f64 ray rounded
// to f32 (the boxes are padded by 4 ulps of the scene's extent, which covers that rounding), and the reference's f64
// primitive tests in the reference's operation order at the leaves.  Distances and barycentrics of a hit are therefore
// bit-identical to the reference's for the same primitive; only exact-t ties can resolve differently (tree order).
// ANY = true: return at the first primitive other than `skip_prim` hit closer than tmax (shadow rays).
template <bool ANY> __device__ __forceinline__ Hit64 trace_mixed(const DevScene &sc, D3v o, D3v d, double tmax, uint32_t skip_prim) {
    Hit64 best;
    best.prim = 0xFFFFFFFFu;
    best.t = tmax; // +inf for closest-hit queries
    best.bx = best.by = 0.0;
    const float inf = __int_as_float(0x7f800000);
    float tbound = ANY ? __double2float_ru(tmax) : inf;
    const float ox = (float)o.x, oy = (float)o.y, oz = (float)o.z, dx = (float)d.x, dy = (float)d.y, dz = (float)d.z;
    const float eps = 8.27180613e-25f;
    const float idx = 1.0f / (fabsf(dx) > eps ? dx : copysignf(eps, dx));
    const float idy = 1.0f / (fabsf(dy) > eps ? dy : copysignf(eps, dy));
    const float idz = 1.0f / (fabsf(dz) > eps ? dz : copysignf(eps, dz));
    const float oix = -ox * idx, oiy = -oy * idy, oiz = -oz * idz;
    int32_t stack[PATHS_FAST_STACK];
    int sp = 0;
    int32_t node = sc.fast_root;
    if (node == (int32_t)0x80000000) return best;
    const float4 *__restrict__ nodes = sc.fast_nodes;
    const float4 *__restrict__ leaves = sc.fast_leaves;
    for (;;) {
        while (node >= 0) {
            const float4 n0 = __ldg(nodes + 4 * node), n1 = __ldg(nodes + 4 * node + 1), n2 = __ldg(nodes + 4 * node + 2), n3 = __ldg(nodes + 4 * node + 3);
            // explicit fma: this translation unit is compiled with -fmad=false for the f64 reference arithmetic
            float a0 = __fmaf_rn(n0.x, idx, oix), a1 = __fmaf_rn(n0.y, idx, oix), a2 = __fmaf_rn(n0.z, idy, oiy), a3 = __fmaf_rn(n0.w, idy, oiy);
            float a4 = __fmaf_rn(n2.x, idz, oiz), a5 = __fmaf_rn(n2.y, idz, oiz);
            float tn0 = fmaxf(fmaxf(fminf(a0, a1), fminf(a2, a3)), fmaxf(fminf(a4, a5), 0.0f));
            float tf0 = fminf(fminf(fmaxf(a0, a1), fmaxf(a2, a3)), fminf(fmaxf(a4, a5), tbound)) * 1.0000004f;
            float b0 = __fmaf_rn(n1.x, idx, oix), b1 = __fmaf_rn(n1.y, idx, oix), b2 = __fmaf_rn(n1.z, idy, oiy), b3 = __fmaf_rn(n1.w, idy, oiy);
            float b4 = __fmaf_rn(n2.z, idz, oiz), b5 = __fmaf_rn(n2.w, idz, oiz);
            float tn1 = fmaxf(fmaxf(fminf(b0, b1), fminf(b2, b3)), fmaxf(fminf(b4, b5), 0.0f));
            float tf1 = fminf(fminf(fmaxf(b0, b1), fmaxf(b2, b3)), fminf(fmaxf(b4, b5), tbound)) * 1.0000004f;
            bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
            int32_t c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (h0 && h1) {
                bool first0 = tn0 <= tn1;
                node = first0 ? c0 : c1;
                stack_push(stack, sp, first0 ? c1 : c0);
            } else if (h0) node = c0;
            else if (h1) node = c1;
            else {
                if (sp == 0) return best;
                node = stack[--sp];
            }
        }
        {
            uint32_t code = (uint32_t)~node;
            uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
            uint32_t tag_next = __float_as_uint(__ldg(leaves + 3 * first + 2).w);
            for (uint32_t k = 0; k < cnt; k++) {
                const uint32_t tag = tag_next;
                if (k + 1 < cnt) tag_next = __float_as_uint(__ldg(leaves + 3 * (first + k + 1) + 2).w); // in flight during this test
                uint32_t prim = tag & 0x7FFFFFFFu;
                if (ANY && prim == skip_prim) continue;
                double t, bx, by;
                if (prim_hit64_tagged(sc, tag, o, d, t, bx, by)) {
                    // src/bvh.rs:93-104: the first hit is taken as it is (even t = inf), later ones only when strictly closer
                    if (ANY ? t < best.t : (best.prim == 0xFFFFFFFFu || t < best.t)) {
                        best.prim = prim;
                        best.t = t;
                        best.bx = bx;
                        best.by = by;
                        if (ANY) return best;
                        tbound = __double2float_ru(t);
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
#define PATHS_CHILD_EMPTY ((int32_t)0x7FFFFFFF)

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
#if PATHS_INNER_SELECT
        // one predicated store and one predicated load instead of a four-way branch: the four outcomes (both / left / right / none)
        // ran one after the other at 4-5 active lanes each (ncu source page: push and pop 7.3 % of the kernel's warp instructions)
        const bool both = h0 && h1, any = h0 || h1, first0 = tn0 <= tn1;
        const int32_t near = (h0 && (!h1 || first0)) ? c0 : c1, far = first0 ? c1 : c0;
        if (both) stack_push(stack, sp, far);
        int32_t next = near;
        if (!any) {
            next = PATHS_TRAV_DONE;
            if (sp > 0) next = stack[--sp];
        }
        node = next;
#else
        if (h0 && h1) {
            bool first0 = tn0 <= tn1;
            node = first0 ? c0 : c1;
            stack_push(stack, sp, first0 ? c1 : c0);
        } else if (h0) node = c0;
        else if (h1) node = c1;
        else node = sp > 0 ? stack[--sp] : PATHS_TRAV_DONE;
#endif
    }
    // one leaf (node < 0, not DONE): tests its primitives, then pops.  Returns true when ANY and something was hit.
    template <bool ANY> __device__ __forceinline__ bool leaf(const float4 *__restrict__ leaves, int32_t *stack, uint32_t skip_prim) {
        if (leaf_test<ANY>(leaves, node, skip_prim)) return true;
        node = sp > 0 ? stack[--sp] : PATHS_TRAV_DONE;
        return false;
    }
    __device__ __forceinline__ int32_t pop(int32_t *stack) { return sp > 0 ? stack[--sp] : PATHS_TRAV_DONE; }
    // tests the primitives of leaf `leaf_code` (< 0) without touching the traversal position
    template <bool ANY> __device__ __forceinline__ bool leaf_test(const float4 *__restrict__ leaves, int32_t leaf_code, uint32_t skip_prim) {
        uint32_t code = (uint32_t)~leaf_code;
        uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
        const RayFrame frame = ray_frame(dx, dy, dz); // recomputed per leaf visit: cheaper than six more live registers
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
                    if (ANY ? tf < t : closer(tf, p, t, prim)) {
                        t = tf; prim = p; bx = by = 0.0f;
                        if (ANY) return true;
                    }
                }
            } else {
                TriHit th;
                if (triangle_hit_wt(s0, s1, s2, frame, ox, oy, oz, dx, dy, dz, t, th) && (ANY ? th.t < t : closer(th.t, p, t, prim))) {
                    t = th.t; prim = p; bx = th.bx; by = th.by;
                    if (ANY) return true;
                }
            }
        }
        return false;
    }
};

// Mixed mode, resumable: trace_mixed cut into the same bounded steps as TravF (f32 slab tests over the production two-wide tree, the
// reference's f64 primitive tests at the leaves, the same visiting order and acceptance rule), so that k_extend_refill_m /
// k_shadow_refill_m can refill lanes.  A ray's answer is bit-identical to trace_mixed's: only the interleaving of lanes changes.
struct TravM {
    D3v o, d;
    float idx, idy, idz, oix, oiy, oiz, tbound;
    double t, bx, by;
    uint32_t prim;
    int32_t node;
    int sp;

    template <bool ANY> __device__ __forceinline__ void init(D3v o_, D3v d_, double tmax, int32_t root) {
        o = o_; d = d_;
        const float ox = (float)o.x, oy = (float)o.y, oz = (float)o.z, dx = (float)d.x, dy = (float)d.y, dz = (float)d.z;
        const float eps = 8.27180613e-25f;
        idx = 1.0f / (fabsf(dx) > eps ? dx : copysignf(eps, dx));
        idy = 1.0f / (fabsf(dy) > eps ? dy : copysignf(eps, dy));
        idz = 1.0f / (fabsf(dz) > eps ? dz : copysignf(eps, dz));
        oix = -ox * idx; oiy = -oy * idy; oiz = -oz * idz;
        tbound = ANY ? __double2float_ru(tmax) : __int_as_float(0x7f800000);
        t = tmax; bx = by = 0.0;
        prim = 0xFFFFFFFFu;
        node = root; // an empty tree's root is PATHS_TRAV_DONE itself
        sp = 0;
    }
    __device__ __forceinline__ void inner(const float4 *__restrict__ nodes, int32_t *stack) {
        const float4 n0 = __ldg(nodes + 4 * node), n1 = __ldg(nodes + 4 * node + 1), n2 = __ldg(nodes + 4 * node + 2), n3 = __ldg(nodes + 4 * node + 3);
        // explicit fma: the f64 translation unit is compiled with -fmad=false
        float a0 = __fmaf_rn(n0.x, idx, oix), a1 = __fmaf_rn(n0.y, idx, oix), a2 = __fmaf_rn(n0.z, idy, oiy), a3 = __fmaf_rn(n0.w, idy, oiy);
        float a4 = __fmaf_rn(n2.x, idz, oiz), a5 = __fmaf_rn(n2.y, idz, oiz);
        float tn0 = fmaxf(fmaxf(fminf(a0, a1), fminf(a2, a3)), fmaxf(fminf(a4, a5), 0.0f));
        float tf0 = fminf(fminf(fmaxf(a0, a1), fmaxf(a2, a3)), fminf(fmaxf(a4, a5), tbound)) * 1.0000004f;
        float b0 = __fmaf_rn(n1.x, idx, oix), b1 = __fmaf_rn(n1.y, idx, oix), b2 = __fmaf_rn(n1.z, idy, oiy), b3 = __fmaf_rn(n1.w, idy, oiy);
        float b4 = __fmaf_rn(n2.z, idz, oiz), b5 = __fmaf_rn(n2.w, idz, oiz);
        float tn1 = fmaxf(fmaxf(fminf(b0, b1), fminf(b2, b3)), fmaxf(fminf(b4, b5), 0.0f));
        float tf1 = fminf(fminf(fmaxf(b0, b1), fmaxf(b2, b3)), fminf(fmaxf(b4, b5), tbound)) * 1.0000004f;
        bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
        int32_t c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
        if (h0 && h1) {
            bool first0 = tn0 <= tn1;
            node = first0 ? c0 : c1;
            stack_push(stack, sp, first0 ? c1 : c0);
        } else if (h0) node = c0;
        else if (h1) node = c1;
        else node = sp > 0 ? stack[--sp] : PATHS_TRAV_DONE;
    }
    // one leaf, then pop.  Returns true when ANY and a primitive other than skip_prim was hit closer than the bound.
    template <bool ANY> __device__ __forceinline__ bool leaf(const DevScene &sc, const float4 *__restrict__ leaves, int32_t *stack, uint32_t skip_prim) {
        if (leaf_test<ANY>(sc, leaves, node, skip_prim)) return true;
        node = sp > 0 ? stack[--sp] : PATHS_TRAV_DONE;
        return false;
    }
    __device__ __forceinline__ int32_t pop(int32_t *stack) { return sp > 0 ? stack[--sp] : PATHS_TRAV_DONE; }
    // tests the primitives of leaf `leaf_code` (< 0) without touching the traversal position.  A leaf may be tested later than it was
    // reached (k_extend2s_m): boxes met meanwhile were tested against a bound that had not shrunk yet, which only adds visits -- every
    // primitive in a box the fresh bound would have culled lies farther than the hit already held, and is turned down here.
    template <bool ANY> __device__ __forceinline__ bool leaf_test(const DevScene &sc, const float4 *__restrict__ leaves, int32_t leaf_code, uint32_t skip_prim) {
        uint32_t code = (uint32_t)~leaf_code;
        uint32_t first = code >> 3, cnt = (code & 7u) + 1u;
        uint32_t tag_next = __float_as_uint(__ldg(leaves + 3 * first + 2).w);
        for (uint32_t k = 0; k < cnt; k++) {
            const uint32_t tag = tag_next;
            if (k + 1 < cnt) tag_next = __float_as_uint(__ldg(leaves + 3 * (first + k + 1) + 2).w); // in flight during this test
            uint32_t p = tag & 0x7FFFFFFFu;
            if (ANY && p == skip_prim) continue;
            double tt, hx, hy;
            if (prim_hit64_tagged(sc, tag, o, d, tt, hx, hy)) {
                // src/bvh.rs:93-104: the first hit is taken as it is (even t = inf), later ones only when strictly closer
                if (ANY ? tt < t : (prim == 0xFFFFFFFFu || tt < t)) {
                    prim = p; t = tt; bx = hx; by = hy;
                    if (ANY) return true;
                    tbound = __double2float_ru(tt);
                }
            }
        }
        return false;
    }
};

// ------------------------------------------------------------------ four-wide traversal, stack in shared memory
// Node = one 128-byte line: six float4 rows {lo.x, hi.x, lo.y, hi.y, lo.z, hi.z} over the four children + their codes.
// Hit children are sorted by entry distance with a 5-exchange network; the nearest is entered, the rest are pushed far-first.
// The stack lives in shared memory as stack[level * blockDim.x + thread] (conflict-free); its depth is sized by the host from
// the tree (3 * depth + 1 entries always suffice), so there is no overflow path.
struct TravF4 : TravF {
    // SORT = false (visibility rays): any occluder will do, so the hit children are taken in slot order and the exchange network
    // (25 of ~110 instructions per node) is left out
    template <bool SORT = true> __device__ __forceinline__ void inner4(const float4 *__restrict__ nodes4, int32_t *sstack, uint32_t stride) {
        const float4 *n = nodes4 + 8 * (size_t)node;
        const float4 lox = __ldg(n), hix = __ldg(n + 1), loy = __ldg(n + 2), hiy = __ldg(n + 3), loz = __ldg(n + 4), hiz = __ldg(n + 5);
        const float4 chf = __ldg(n + 6);
        const float inf = __int_as_float(0x7f800000);
        float k0, k1, k2, k3;
#define PATHS_BOX4(c, key)                                                                                                  \
    {                                                                                                                       \
        float x0 = lox.c * idx - oix, x1 = hix.c * idx - oix, y0 = loy.c * idy - oiy, y1 = hiy.c * idy - oiy;               \
        float z0 = loz.c * idz - oiz, z1 = hiz.c * idz - oiz;                                                               \
        float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));                                  \
        float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), t)) * 1.0000004f;                        \
        key = tn <= tf ? tn : inf;                                                                                          \
    }
        PATHS_BOX4(x, k0)
        PATHS_BOX4(y, k1)
        PATHS_BOX4(z, k2)
        PATHS_BOX4(w, k3)
#undef PATHS_BOX4
        int32_t c0 = __float_as_int(chf.x), c1 = __float_as_int(chf.y), c2 = __float_as_int(chf.z), c3 = __float_as_int(chf.w);
        // unused slots (an inverted box still passes a min/max slab test) are marked by their code
        k0 = c0 == PATHS_CHILD_EMPTY ? inf : k0;
        k1 = c1 == PATHS_CHILD_EMPTY ? inf : k1;
        k2 = c2 == PATHS_CHILD_EMPTY ? inf : k2;
        k3 = c3 == PATHS_CHILD_EMPTY ? inf : k3;
#define PATHS_CE(ka, ca, kb, cb)          \
    if (kb < ka) {                        \
        float tk = ka; ka = kb; kb = tk;  \
        int32_t tc = ca; ca = cb; cb = tc;\
    }
        if (SORT) {
            PATHS_CE(k0, c0, k1, c1)
            PATHS_CE(k2, c2, k3, c3)
            PATHS_CE(k0, c0, k2, c2)
            PATHS_CE(k1, c1, k3, c3)
            PATHS_CE(k1, c1, k2, c2)
            if (k3 < inf) sstack[(sp++) * stride] = c3;
            if (k2 < inf) sstack[(sp++) * stride] = c2;
            if (k1 < inf) sstack[(sp++) * stride] = c1;
            if (k0 < inf) node = c0;
            else node = sp > 0 ? sstack[(--sp) * stride] : PATHS_TRAV_DONE;
        } else {
            // unsorted: hits are not a prefix any more; enter the last hit child, push the others
            int32_t next = PATHS_TRAV_DONE;
            if (k0 < inf) next = c0;
            if (k1 < inf) { if (next != PATHS_TRAV_DONE) sstack[(sp++) * stride] = next; next = c1; }
            if (k2 < inf) { if (next != PATHS_TRAV_DONE) sstack[(sp++) * stride] = next; next = c2; }
            if (k3 < inf) { if (next != PATHS_TRAV_DONE) sstack[(sp++) * stride] = next; next = c3; }
            node = next != PATHS_TRAV_DONE ? next : (sp > 0 ? sstack[(--sp) * stride] : PATHS_TRAV_DONE);
        }
#undef PATHS_CE
    }
    __device__ __forceinline__ int32_t pop4(const int32_t *sstack, uint32_t stride) { return sp > 0 ? sstack[(--sp) * stride] : PATHS_TRAV_DONE; }
};


// Mixed mode over the four-wide nodes (scenes above 64k primitives): the slab tests of TravF4 against TravM's f32 bound, the f64
// primitive tests of TravM at the leaves, stack in shared memory.  Half as many dependent node fetches per ray as the two-wide walk,
// which is what a kernel that waits on memory wants.  Visiting order differs from trace_mixed's, so exact-t ties may resolve to
// another primitive; the closest distance does not depend on the tree.
struct TravM4 : TravM {
    template <bool SORT = true> __device__ __forceinline__ void inner4(const float4 *__restrict__ nodes4, int32_t *sstack, uint32_t stride) {
        const float4 *n = nodes4 + 8 * (size_t)node;
        const float4 lox = __ldg(n), hix = __ldg(n + 1), loy = __ldg(n + 2), hiy = __ldg(n + 3), loz = __ldg(n + 4), hiz = __ldg(n + 5);
        const float4 chf = __ldg(n + 6);
        const float inf = __int_as_float(0x7f800000);
        float k0, k1, k2, k3;
#define PATHS_BOX4M(c, key)                                                                                                 \
    {                                                                                                                       \
        float x0 = __fmaf_rn(lox.c, idx, oix), x1 = __fmaf_rn(hix.c, idx, oix), y0 = __fmaf_rn(loy.c, idy, oiy);            \
        float y1 = __fmaf_rn(hiy.c, idy, oiy), z0 = __fmaf_rn(loz.c, idz, oiz), z1 = __fmaf_rn(hiz.c, idz, oiz);            \
        float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));                                  \
        float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tbound)) * 1.0000004f;                   \
        key = tn <= tf ? tn : inf;                                                                                          \
    }
        PATHS_BOX4M(x, k0)
        PATHS_BOX4M(y, k1)
        PATHS_BOX4M(z, k2)
        PATHS_BOX4M(w, k3)
#undef PATHS_BOX4M
        int32_t c0 = __float_as_int(chf.x), c1 = __float_as_int(chf.y), c2 = __float_as_int(chf.z), c3 = __float_as_int(chf.w);
        k0 = c0 == PATHS_CHILD_EMPTY ? inf : k0;
        k1 = c1 == PATHS_CHILD_EMPTY ? inf : k1;
        k2 = c2 == PATHS_CHILD_EMPTY ? inf : k2;
        k3 = c3 == PATHS_CHILD_EMPTY ? inf : k3;
#define PATHS_CE(ka, ca, kb, cb)          \
    if (kb < ka) {                        \
        float tk = ka; ka = kb; kb = tk;  \
        int32_t tc = ca; ca = cb; cb = tc;\
    }
        if (SORT) {
            PATHS_CE(k0, c0, k1, c1)
            PATHS_CE(k2, c2, k3, c3)
            PATHS_CE(k0, c0, k2, c2)
            PATHS_CE(k1, c1, k3, c3)
            PATHS_CE(k1, c1, k2, c2)
            if (k3 < inf) sstack[(sp++) * stride] = c3;
            if (k2 < inf) sstack[(sp++) * stride] = c2;
            if (k1 < inf) sstack[(sp++) * stride] = c1;
            if (k0 < inf) node = c0;
            else node = sp > 0 ? sstack[(--sp) * stride] : PATHS_TRAV_DONE;
        } else {
            int32_t next = PATHS_TRAV_DONE;
            if (k0 < inf) next = c0;
            if (k1 < inf) { if (next != PATHS_TRAV_DONE) sstack[(sp++) * stride] = next; next = c1; }
            if (k2 < inf) { if (next != PATHS_TRAV_DONE) sstack[(sp++) * stride] = next; next = c2; }
            if (k3 < inf) { if (next != PATHS_TRAV_DONE) sstack[(sp++) * stride] = next; next = c3; }
            node = next != PATHS_TRAV_DONE ? next : (sp > 0 ? sstack[(--sp) * stride] : PATHS_TRAV_DONE);
        }
#undef PATHS_CE
    }
    __device__ __forceinline__ int32_t pop4(const int32_t *sstack, uint32_t stride) { return sp > 0 ? sstack[(--sp) * stride] : PATHS_TRAV_DONE; }
};

// plain per-lane closest hit over the four-wide nodes (camera segment of large scenes, and the per-ray seam that tests it)
__device__ __forceinline__ HitF trace_plain4(const DevScene &sc, float ox, float oy, float oz, float dx, float dy, float dz, int32_t *sstack,
                                             uint32_t stride) {
    TravF4 tv;
    tv.init(ox, oy, oz, dx, dy, dz, __int_as_float(0x7f800000), sc.fast4_root);
    while (tv.node != PATHS_TRAV_DONE) {
        if (tv.node >= 0) tv.inner4(sc.fast4_nodes, sstack, stride);
        else {
            tv.leaf_test<false>(sc.fast_leaves, tv.node, 0xFFFFFFFFu);
            tv.node = tv.pop4(sstack, stride);
        }
    }
    HitF h;
    h.prim = tv.prim; h.t = tv.t; h.bx = tv.bx; h.by = tv.by;
    return h;
}

} // namespace paths
