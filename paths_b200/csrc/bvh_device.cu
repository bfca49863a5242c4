// bvh_device.cu -- the production acceleration structure built ON the GPU: replaces the host builders
// (src/bvh.rs:160-343 in the reference, bvh_fast.cpp here) for scenes that change or are large.
//
//   k_prim_bounds   conservative f32 box per primitive from the f64 prim table already in HBM; hidden primitives
//                   (flat own box, never hit by the reference: src/bvh.rs:16) are dropped like in bvh_fast.cpp
//   k_classify      "giant" primitives (the scenes' 1e6-radius plane spheres) are kept out of the morton domain -- they
//                   are what collapses the reference's own morton scale (src/bvh.rs:178-182) -- and chained above the root
//   k_morton        63-bit morton key of the box centre inside the bounds of the ordinary primitives
//   cub radix sort  (key, prim) pairs
//   k_hierarchy     Karras 2012: one thread per inner node finds its range and split from common key prefixes
//   k_fit           bottom-up boxes through per-node arrival counters, with the SAH decision to close a subtree of <= 4
//                   primitives into one leaf (same cost constants as bvh_fast.cpp)
//   k_emit2 / k_emit_leaves / k_emit4_level   the two node layouts and the leaf stream the traversal kernels read
//
// Closest hit does not depend on the shape of the tree (only exact-t ties do), so a morton tree gives the same images as the
// SAH tree; it is built in milliseconds instead of a second and traverses somewhat slower (numbers in DESIGN.md).
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <limits>

#include "scene_device.h"

namespace paths {
namespace {

#define BD_CU(call)                                                                         \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess) {                                                            \
            err = std::string(#call) + ": " + cudaGetErrorString(e_);                       \
            return false;                                                                   \
        }                                                                                   \
    } while (0)

const float kTraversalCost = 1.0f, kPrimCost = 1.2f; // bvh_fast.cpp
const uint32_t kMaxLeaf = 4;
const uint32_t kMaxGiants = 48;

struct BuildCtl {                 // device-side scalars of one build
    unsigned long long log_sum;   // sum of floor(256 * (log2(extent) + 200)) over visible prims
    uint32_t n_visible;
    uint32_t n_normal;            // visible and not giant
    uint32_t n_giants;            // may exceed kMaxGiants: then nothing is treated as giant
    uint32_t giants[kMaxGiants];
    int32_t cmin[3], cmax[3];     // ordered-int bounds of the ordinary primitives' box centres
    uint32_t n_nodes4;            // four-wide nodes allocated so far
    uint32_t qcount[2];           // emit4 level queues
    uint32_t levels4;
    uint32_t depth2;
};

__device__ __forceinline__ int32_t float_ordered(float f) {
    int32_t i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __host__ __forceinline__ float ordered_float(int32_t i) {
    int32_t b = i >= 0 ? i : i ^ 0x7FFFFFFF;
    float f;
    memcpy(&f, &b, 4);
    return f;
}

// bvh_fast.cpp build_fast_tree(): outward-rounded box, covering the f32-rounded vertices the f32 kernels intersect, padded
__global__ void k_prim_bounds(const uint4 *__restrict__ meta, const double *__restrict__ geom, uint32_t n, float pad, float4 *blo, float4 *bhi,
                              BuildCtl *ctl) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *g = geom + (size_t)i * 12;
    bool tri = (meta[i].x & PATHS_PM_TRI) != 0;
    double lo[3], hi[3];
    bool hidden = false;
    if (tri) {
        for (int a = 0; a < 3; a++) {
            lo[a] = fmin(g[a], fmin(g[3 + a], g[6 + a]));
            hi[a] = fmax(g[a], fmax(g[3 + a], g[6 + a]));
            hidden = hidden || lo[a] == hi[a];
        }
    } else {
        for (int a = 0; a < 3; a++) {
            lo[a] = g[a] - g[3];
            hi[a] = g[a] + g[3];
        }
    }
    float flo[3], fhi[3];
    for (int a = 0; a < 3; a++) {
        flo[a] = __double2float_rd(lo[a]);
        fhi[a] = __double2float_ru(hi[a]);
        if (tri) {
            for (int k = 0; k < 3; k++) {
                float v = (float)g[3 * k + a];
                flo[a] = fminf(flo[a], v);
                fhi[a] = fmaxf(fhi[a], v);
            }
            flo[a] -= pad;
            fhi[a] += pad;
        }
    }
    float ext = fmaxf(fhi[0] - flo[0], fmaxf(fhi[1] - flo[1], fhi[2] - flo[2]));
    blo[i] = make_float4(flo[0], flo[1], flo[2], hidden ? -1.0f : ext);
    bhi[i] = make_float4(fhi[0], fhi[1], fhi[2], 0.0f);
    if (!hidden) {
        float l2 = log2f(fmaxf(ext, 1e-30f)) + 200.0f; // > 0 for every finite extent
        atomicAdd(&ctl->log_sum, (unsigned long long)(l2 * 256.0f));
        atomicAdd(&ctl->n_visible, 1u);
    }
}

// giant: extent above 1024 x the geometric mean extent of the scene's visible primitives
__device__ __forceinline__ float giant_threshold(const BuildCtl *ctl) {
    float mean_l2 = (float)((double)ctl->log_sum / (256.0 * (double)max(ctl->n_visible, 1u))) - 200.0f;
    return exp2f(mean_l2 + 10.0f);
}

__global__ void k_classify(uint32_t n, float4 *blo, const float4 *__restrict__ bhi, BuildCtl *ctl) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 lo = blo[i], hi = bhi[i];
    if (lo.w < 0.0f) return; // hidden
    if (lo.w > giant_threshold(ctl)) {
        uint32_t g = atomicAdd(&ctl->n_giants, 1u);
        if (g < kMaxGiants) ctl->giants[g] = i;
        blo[i].w = -2.0f; // provisional; k_morton undoes it when there are too many giants
        return;
    }
    float c[3] = {0.5f * lo.x + 0.5f * hi.x, 0.5f * lo.y + 0.5f * hi.y, 0.5f * lo.z + 0.5f * hi.z};
    for (int a = 0; a < 3; a++) {
        atomicMin(&ctl->cmin[a], float_ordered(c[a]));
        atomicMax(&ctl->cmax[a], float_ordered(c[a]));
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long x) { // 21 bits -> every third bit
    x &= 0x1FFFFFull;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// keys: ordinary prims get their morton code; hidden prims and giants sort to the end (all ones)
__global__ void k_morton(uint32_t n, const float4 *__restrict__ blo, const float4 *__restrict__ bhi, const BuildCtl *ctl, int giants_ok,
                         unsigned long long *keys, uint32_t *vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 lo = blo[i], hi = bhi[i];
    vals[i] = i;
    bool ordinary = lo.w >= 0.0f || (lo.w == -2.0f && !giants_ok);
    if (!ordinary) {
        keys[i] = ~0ull;
        return;
    }
    float c[3] = {0.5f * lo.x + 0.5f * hi.x, 0.5f * lo.y + 0.5f * hi.y, 0.5f * lo.z + 0.5f * hi.z};
    unsigned long long q[3];
    for (int a = 0; a < 3; a++) {
        float mn = ordered_float(ctl->cmin[a]), mx = ordered_float(ctl->cmax[a]);
        float s = mx > mn ? (c[a] - mn) / (mx - mn) : 0.0f;
        s = fminf(fmaxf(s, 0.0f), 1.0f); // giants that were put back (too many of them) clamp to the border
        q[a] = (unsigned long long)(s * 2097151.0f);
    }
    keys[i] = spread21(q[0]) | (spread21(q[1]) << 1) | (spread21(q[2]) << 2);
}

// ------------------------------------------------------------------ hierarchy (Karras 2012, "Maximizing parallelism in the construction of BVHs")
// Inner node i covers the sorted range [first[i], last[i]]; children are inner nodes (>= 0) or sorted positions (~pos < 0).
struct Topo {
    int32_t *left, *right;   // [n - 1]
    int32_t *parent;         // [2n - 1]: inner nodes 0..n-2, then leaves n-1..2n-2
    uint32_t *first, *last;  // [n - 1]
};

__device__ __forceinline__ int delta(const unsigned long long *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j); // equal keys: the position breaks the tie
    return __clzll((long long)(a ^ b));
}

__global__ void k_hierarchy(const unsigned long long *__restrict__ keys, int n, Topo tp) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int32_t lc = lo == gamma ? ~gamma : gamma;
    int32_t rc = hi == gamma + 1 ? ~(gamma + 1) : gamma + 1;
    tp.left[i] = lc;
    tp.right[i] = rc;
    tp.first[i] = (uint32_t)lo;
    tp.last[i] = (uint32_t)hi;
    tp.parent[lc >= 0 ? lc : (n - 1) + ~lc] = i;
    tp.parent[rc >= 0 ? rc : (n - 1) + ~rc] = i;
    if (i == 0) tp.parent[0] = -1;
}

// ------------------------------------------------------------------ boxes, SAH leaf decision, depth: bottom-up
struct Fit {
    float4 *nlo, *nhi;     // [n - 1] inner-node boxes; nlo.w = subtree SAH cost, nhi.w = bits: closed (1) | depth << 1
    uint32_t *arrived;     // [n - 1]
};

__device__ __forceinline__ float half_area(float4 lo, float4 hi) {
    float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_fit(int n, const uint32_t *__restrict__ sorted, const float4 *__restrict__ blo, const float4 *__restrict__ bhi, Topo tp, Fit ft,
                      BuildCtl *ctl) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int node = tp.parent[(n - 1) + k];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&ft.arrived[node], 1u) == 0u) return; // the sibling subtree is not finished: its thread carries on
        __threadfence();
        float4 lo[2], hi[2];
        float cost[2];
        uint32_t depth[2];
        int32_t ch[2] = {tp.left[node], tp.right[node]};
        for (int c = 0; c < 2; c++) {
            if (ch[c] < 0) {
                uint32_t prim = sorted[~ch[c]];
                lo[c] = blo[prim];
                hi[c] = bhi[prim];
                cost[c] = kPrimCost;
                depth[c] = 1;
            } else {
                lo[c] = __ldcg(ft.nlo + ch[c]);
                hi[c] = __ldcg(ft.nhi + ch[c]);
                cost[c] = lo[c].w;
                depth[c] = __float_as_uint(hi[c].w) >> 1;
            }
        }
        float4 l = make_float4(fminf(lo[0].x, lo[1].x), fminf(lo[0].y, lo[1].y), fminf(lo[0].z, lo[1].z), 0.0f);
        float4 h = make_float4(fmaxf(hi[0].x, hi[1].x), fmaxf(hi[0].y, hi[1].y), fmaxf(hi[0].z, hi[1].z), 0.0f);
        uint32_t count = tp.last[node] - tp.first[node] + 1;
        float area = fmaxf(half_area(l, h), 1e-30f);
        float split = kTraversalCost + (half_area(lo[0], hi[0]) * cost[0] + half_area(lo[1], hi[1]) * cost[1]) / area;
        float leaf = kPrimCost * (float)count;
        bool closed = count <= kMaxLeaf && leaf <= split;
        uint32_t dep = closed ? 1u : 1u + max(depth[0], depth[1]);
        l.w = closed ? leaf : split;
        h.w = __uint_as_float((dep << 1) | (closed ? 1u : 0u));
        ft.nlo[node] = l;
        ft.nhi[node] = h;
        if (node == 0) ctl->depth2 = dep;
        node = tp.parent[node];
    }
}

// ------------------------------------------------------------------ emission
__device__ __forceinline__ int32_t leaf_code(uint32_t first, uint32_t count) { return ~(int32_t)((first << 3) | (count - 1)); }

// code of child `c` of an inner node, with its box: a sorted position, a closed subtree (one leaf) or an inner node
__device__ __forceinline__ int32_t child_code(int32_t c, const uint32_t *__restrict__ sorted, const float4 *__restrict__ blo,
                                              const float4 *__restrict__ bhi, const Topo &tp, const Fit &ft, float4 &lo, float4 &hi) {
    if (c < 0) {
        uint32_t pos = (uint32_t)~c, prim = sorted[pos];
        lo = blo[prim];
        hi = bhi[prim];
        return leaf_code(pos, 1);
    }
    lo = ft.nlo[c];
    hi = ft.nhi[c];
    if (__float_as_uint(hi.w) & 1u) return leaf_code(tp.first[c], tp.last[c] - tp.first[c] + 1);
    return c;
}

// two-wide nodes, index = Karras inner-node index (nodes under a closed subtree are written but never referenced)
__global__ void k_emit2(int n, const uint32_t *__restrict__ sorted, const float4 *__restrict__ blo, const float4 *__restrict__ bhi, Topo tp, Fit ft,
                        float4 *nodes) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    float4 l0, h0, l1, h1;
    int32_t c0 = child_code(tp.left[i], sorted, blo, bhi, tp, ft, l0, h0);
    int32_t c1 = child_code(tp.right[i], sorted, blo, bhi, tp, ft, l1, h1);
    nodes[4 * (size_t)i + 0] = make_float4(l0.x, h0.x, l0.y, h0.y);
    nodes[4 * (size_t)i + 1] = make_float4(l1.x, h1.x, l1.y, h1.y);
    nodes[4 * (size_t)i + 2] = make_float4(l0.z, h0.z, l1.z, h1.z);
    nodes[4 * (size_t)i + 3] = make_float4(__int_as_float(c0), __int_as_float(c1), 0.0f, 0.0f);
}

// leaf stream in sorted order (layout: scene_device.cu), giants appended behind the ordinary primitives
__global__ void k_emit_leaves(uint32_t n_slots, uint32_t n_normal, const uint32_t *__restrict__ sorted, const BuildCtl *ctl,
                              const uint4 *__restrict__ meta, const double *__restrict__ geom, float4 *leaves) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_slots) return;
    uint32_t prim = k < n_normal ? sorted[k] : ctl->giants[k - n_normal];
    const double *g = geom + (size_t)prim * 12;
    float4 *s = leaves + 3 * (size_t)k;
    if (meta[prim].x & PATHS_PM_TRI) {
        s[0] = make_float4((float)g[0], (float)g[1], (float)g[2], 0.0f);
        s[1] = make_float4((float)g[3], (float)g[4], (float)g[5], 0.0f);
        s[2] = make_float4((float)g[6], (float)g[7], (float)g[8], __uint_as_float(prim));
    } else {
        s[0] = make_float4(__int_as_float(__double2loint(g[0])), __int_as_float(__double2hiint(g[0])), __int_as_float(__double2loint(g[1])),
                           __int_as_float(__double2hiint(g[1])));
        s[1] = make_float4(__int_as_float(__double2loint(g[2])), __int_as_float(__double2hiint(g[2])), __int_as_float(__double2loint(g[3])),
                           __int_as_float(__double2hiint(g[3])));
        s[2] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(prim | 0x80000000u));
    }
}

// four-wide nodes, one tree level per launch: entry = (binary inner node, index of the four-wide node it becomes).  The node's
// children are its binary children with the largest open one replaced by its own two until four slots are used (bvh_fast.cpp emit4).
__global__ void k_emit4_level(const uint2 *__restrict__ qin, uint2 *qout, int which, uint32_t level, const uint32_t *__restrict__ sorted,
                              const float4 *__restrict__ blo, const float4 *__restrict__ bhi, Topo tp, Fit ft, float4 *nodes4, BuildCtl *ctl) {
    const uint32_t count = ctl->qcount[which];
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < count; e += gridDim.x * blockDim.x) {
        if (e == 0) ctl->levels4 = level + 1;
        uint2 ent = qin[e];
        int32_t kids[4];
        int nk = 2;
        kids[0] = tp.left[ent.x];
        kids[1] = tp.right[ent.x];
        while (nk < 4) {
            int best = -1;
            float best_area = -1.0f;
            for (int i = 0; i < nk; i++) {
                if (kids[i] < 0) continue;
                float4 lo = ft.nlo[kids[i]], hi = ft.nhi[kids[i]];
                if (__float_as_uint(hi.w) & 1u) continue; // closed: a leaf
                float a = half_area(lo, hi);
                if (a > best_area) {
                    best_area = a;
                    best = i;
                }
            }
            if (best < 0) break;
            int32_t open = kids[best];
            kids[best] = tp.left[open];
            kids[nk++] = tp.right[open];
        }
        float lox[4], hix[4], loy[4], hiy[4], loz[4], hiz[4];
        int32_t code[4];
        const float inf = __int_as_float(0x7f800000);
        for (int i = 0; i < 4; i++) {
            if (i < nk) {
                float4 lo, hi;
                int32_t c = child_code(kids[i], sorted, blo, bhi, tp, ft, lo, hi);
                if (c >= 0) { // an inner node: allocate its four-wide node and queue it for the next level
                    uint32_t idx = atomicAdd(&ctl->n_nodes4, 1u);
                    uint32_t q = atomicAdd(&ctl->qcount[which ^ 1], 1u);
                    qout[q] = make_uint2((uint32_t)c, idx);
                    c = (int32_t)idx;
                }
                lox[i] = lo.x; hix[i] = hi.x; loy[i] = lo.y; hiy[i] = hi.y; loz[i] = lo.z; hiz[i] = hi.z;
                code[i] = c;
            } else {
                lox[i] = loy[i] = loz[i] = inf;
                hix[i] = hiy[i] = hiz[i] = -inf;
                code[i] = 0x7FFFFFFF; // PATHS_CHILD_EMPTY
            }
        }
        float4 *o = nodes4 + 8 * (size_t)ent.y;
        o[0] = make_float4(lox[0], lox[1], lox[2], lox[3]);
        o[1] = make_float4(hix[0], hix[1], hix[2], hix[3]);
        o[2] = make_float4(loy[0], loy[1], loy[2], loy[3]);
        o[3] = make_float4(hiy[0], hiy[1], hiy[2], hiy[3]);
        o[4] = make_float4(loz[0], loz[1], loz[2], loz[3]);
        o[5] = make_float4(hiz[0], hiz[1], hiz[2], hiz[3]);
        o[6] = make_float4(__int_as_float(code[0]), __int_as_float(code[1]), __int_as_float(code[2]), __int_as_float(code[3]));
        o[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}
__global__ void k_queue_flip(BuildCtl *ctl, int which) { ctl->qcount[which] = 0; }

} // namespace

// ------------------------------------------------------------------ driver
namespace {
struct Scratch { // frees everything on scope exit
    std::vector<void *> ptrs;
    ~Scratch() {
        for (void *p : ptrs) cudaFree(p);
    }
    template <typename T> bool get(T **out, size_t count, std::string &err) {
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
        if (e != cudaSuccess) {
            err = std::string("cudaMalloc (bvh build): ") + cudaGetErrorString(e);
            return false;
        }
        ptrs.push_back(p);
        *out = (T *)p;
        return true;
    }
};
struct HBox {
    float lo[3], hi[3];
    void grow(const HBox &b) {
        for (int a = 0; a < 3; a++) {
            lo[a] = std::min(lo[a], b.lo[a]);
            hi[a] = std::max(hi[a], b.hi[a]);
        }
    }
};
template <typename T> bool keep(SceneDevice *sd, T **out, size_t count, std::string &err) {
    void *p = nullptr;
    size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
        err = std::string("cudaMalloc (bvh): ") + cudaGetErrorString(e);
        return false;
    }
    sd->allocs.push_back(p);
    sd->bytes += bytes;
    *out = (T *)p;
    return true;
}
} // namespace

// Builds the production tree of the staged scene `sd` on its own GPU and points sd->view at it.  Returns false with
// err == "small" when the scene has too few ordinary primitives for a morton build (the caller then uses the host builder).
bool build_device_tree(SceneDevice *sd, const HostScene &hs, DeviceTreeInfo &info, std::string &err) {
    BD_CU(cudaSetDevice(sd->device));
    const uint32_t n = sd->view.n_prims;
    info = DeviceTreeInfo();
    if (n < 64) {
        err = "small";
        return false;
    }
    cudaEvent_t e0, e1;
    BD_CU(cudaEventCreate(&e0));
    BD_CU(cudaEventCreate(&e1));
    BD_CU(cudaEventRecord(e0, 0));
    Scratch sc;
    BuildCtl *ctl;
    float4 *blo, *bhi;
    unsigned long long *keys, *keys2;
    uint32_t *vals, *vals2;
    if (!sc.get(&ctl, 1, err) || !sc.get(&blo, n, err) || !sc.get(&bhi, n, err) || !sc.get(&keys, n, err) || !sc.get(&keys2, n, err) ||
        !sc.get(&vals, n, err) || !sc.get(&vals2, n, err))
        return false;
    BuildCtl h;
    std::memset(&h, 0, sizeof(h));
    for (int a = 0; a < 3; a++) {
        h.cmin[a] = 0x7FFFFFFF;
        h.cmax[a] = (int32_t)0x80000000;
    }
    BD_CU(cudaMemcpy(ctl, &h, sizeof(h), cudaMemcpyHostToDevice));
    double ext = 1.0; // bvh_fast.cpp: boxes grow by a few ulps of the scene's extent
    if (hs.bounds_lo.x <= hs.bounds_hi.x)
        ext = std::max({std::fabs(hs.bounds_lo.x), std::fabs(hs.bounds_hi.x), std::fabs(hs.bounds_lo.y), std::fabs(hs.bounds_hi.y),
                        std::fabs(hs.bounds_lo.z), std::fabs(hs.bounds_hi.z), 1.0});
    const float pad = (float)(ext * 4.76837158203125e-7);
    const uint32_t T = 256, G = (n + T - 1) / T;
    k_prim_bounds<<<G, T>>>(sd->view.prim_meta, sd->view.prim_geom, n, pad, blo, bhi, ctl);
    k_classify<<<G, T>>>(n, blo, bhi, ctl);
    BD_CU(cudaMemcpy(&h, ctl, sizeof(h), cudaMemcpyDeviceToHost));
    const bool giants_ok = h.n_giants <= kMaxGiants;
    const uint32_t n_giants = giants_ok ? h.n_giants : 0;
    const uint32_t nn = h.n_visible - n_giants; // ordinary primitives: the morton tree's leaves
    if (nn < 64) {
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        err = "small";
        return false;
    }
    k_morton<<<G, T>>>(n, blo, bhi, ctl, giants_ok ? 1 : 0, keys, vals);
    {
        size_t tmp_bytes = 0;
        BD_CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, vals, vals2, (int)n, 0, 64));
        void *tmp;
        if (!sc.get((char **)&tmp, tmp_bytes, err)) return false;
        BD_CU(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, vals, vals2, (int)n, 0, 64));
    }
    const uint32_t *sorted = vals2; // positions [0, nn) hold the ordinary primitives in morton order
    Topo tp;
    Fit ft;
    if (!sc.get(&tp.left, nn - 1, err) || !sc.get(&tp.right, nn - 1, err) || !sc.get(&tp.parent, 2 * (size_t)nn - 1, err) ||
        !sc.get(&tp.first, nn - 1, err) || !sc.get(&tp.last, nn - 1, err) || !sc.get(&ft.nlo, nn - 1, err) || !sc.get(&ft.nhi, nn - 1, err) ||
        !sc.get(&ft.arrived, nn - 1, err))
        return false;
    BD_CU(cudaMemset(ft.arrived, 0, (size_t)(nn - 1) * 4));
    k_hierarchy<<<(nn + T - 1) / T, T>>>(keys2, (int)nn, tp);
    k_fit<<<(nn + T - 1) / T, T>>>((int)nn, sorted, blo, bhi, tp, ft, ctl);
    // ---- two-wide nodes + leaf stream (kept)
    const uint32_t n_chain2 = n_giants;
    const uint32_t n_slots = nn + n_giants;
    float4 *nodes2, *leaves;
    if (!keep(sd, &nodes2, 4 * ((size_t)(nn - 1) + n_chain2), err) || !keep(sd, &leaves, 3 * (size_t)n_slots, err)) return false;
    k_emit2<<<(nn + T - 1) / T, T>>>((int)nn, sorted, blo, bhi, tp, ft, nodes2);
    k_emit_leaves<<<(n_slots + T - 1) / T, T>>>(n_slots, nn, sorted, ctl, sd->view.prim_meta, sd->view.prim_geom, leaves);
    // ---- four-wide nodes, level by level into a scratch array of worst-case size
    const uint32_t n_chain4 = (n_giants + 2) / 3;
    float4 *nodes4_tmp;
    uint2 *q[2];
    if (!sc.get(&nodes4_tmp, 8 * (size_t)(nn - 1), err) || !sc.get(&q[0], nn, err) || !sc.get(&q[1], nn, err)) return false;
    {
        uint2 root_entry = make_uint2(0u, 0u);
        BD_CU(cudaMemcpy(q[0], &root_entry, sizeof(root_entry), cudaMemcpyHostToDevice));
        uint32_t init[4] = {1u, 1u, 0u, 0u}; // n_nodes4 = 1 (the root), qcount = {1, 0}, levels4 = 0
        BD_CU(cudaMemcpy(&ctl->n_nodes4, init, sizeof(init), cudaMemcpyHostToDevice));
    }
    uint32_t level = 0;
    for (bool done = false; !done;) {
        for (int k = 0; k < 16; k++, level++) {
            int w = (int)(level & 1u);
            k_queue_flip<<<1, 1>>>(ctl, w ^ 1);
            k_emit4_level<<<592, 256>>>(q[w], q[w ^ 1], w, level, sorted, blo, bhi, tp, ft, nodes4_tmp, ctl);
        }
        BD_CU(cudaMemcpy(&h, ctl, sizeof(h), cudaMemcpyDeviceToHost));
        done = h.qcount[level & 1u] == 0;
        if (level > 4096) {
            err = "four-wide emission did not terminate";
            return false;
        }
    }
    // the root may itself be closed (<= 4 prims) only when nn <= 4, excluded above: node 0 is always an inner node
    const uint32_t n4 = h.n_nodes4;
    float4 *nodes4;
    if (!keep(sd, &nodes4, 8 * ((size_t)n4 + n_chain4), err)) return false;
    BD_CU(cudaMemcpy(nodes4, nodes4_tmp, (size_t)n4 * 128, cudaMemcpyDeviceToDevice));
    int32_t root2 = 0, root4 = 0;
    uint32_t depth2 = h.depth2, depth4 = h.levels4;
    if (depth2 + n_giants > 60) { // the two-wide kernels keep a 64-entry stack (PATHS_FAST_STACK): a morton tree this deep (long runs
        cudaEventDestroy(e0);     // of equal keys) goes back to the host SAH builder, whose median fallback keeps trees shallow
        cudaEventDestroy(e1);
        err = "small";
        return false;
    }
    // ---- giants: a chain above the morton tree's root, largest first is not needed -- every ray tests them all anyway
    if (n_giants > 0) {
        std::vector<HBox> gb(n_giants);
        for (uint32_t g = 0; g < n_giants; g++) {
            float4 lo, hi;
            BD_CU(cudaMemcpy(&lo, blo + h.giants[g], sizeof(lo), cudaMemcpyDeviceToHost));
            BD_CU(cudaMemcpy(&hi, bhi + h.giants[g], sizeof(hi), cudaMemcpyDeviceToHost));
            gb[g] = HBox{{lo.x, lo.y, lo.z}, {hi.x, hi.y, hi.z}};
        }
        float4 rlo, rhi;
        BD_CU(cudaMemcpy(&rlo, ft.nlo, sizeof(rlo), cudaMemcpyDeviceToHost));
        BD_CU(cudaMemcpy(&rhi, ft.nhi, sizeof(rhi), cudaMemcpyDeviceToHost));
        HBox rest{{rlo.x, rlo.y, rlo.z}, {rhi.x, rhi.y, rhi.z}};
        // two-wide chain: node j = {giant j, everything after it}; built back to front
        std::vector<float4> chain2(4 * (size_t)n_giants);
        int32_t below = 0; // the morton root
        HBox below_box = rest;
        for (int g = (int)n_giants - 1; g >= 0; g--) {
            int32_t code = ~(int32_t)(((nn + (uint32_t)g) << 3) | 0u);
            float4 *c = &chain2[4 * (size_t)g];
            c[0] = make_float4(gb[g].lo[0], gb[g].hi[0], gb[g].lo[1], gb[g].hi[1]);
            c[1] = make_float4(below_box.lo[0], below_box.hi[0], below_box.lo[1], below_box.hi[1]);
            c[2] = make_float4(gb[g].lo[2], gb[g].hi[2], below_box.lo[2], below_box.hi[2]);
            float4 k = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            std::memcpy(&k.x, &code, 4);
            std::memcpy(&k.y, &below, 4);
            c[3] = k;
            below = (int32_t)(nn - 1) + g;
            below_box.grow(gb[g]);
        }
        root2 = below;
        depth2 += n_giants;
        BD_CU(cudaMemcpy(nodes2 + 4 * (size_t)(nn - 1), chain2.data(), chain2.size() * sizeof(float4), cudaMemcpyHostToDevice));
        // four-wide chain: node j = {up to three giants, everything after them}
        std::vector<float> chain4(32 * (size_t)n_chain4);
        below = 0;
        below_box = rest;
        for (int j = (int)n_chain4 - 1; j >= 0; j--) {
            float *o = &chain4[32 * (size_t)j];
            int32_t codes[4];
            HBox boxes[4];
            int nk = 0;
            for (uint32_t g = 3 * (uint32_t)j; g < std::min(n_giants, 3 * (uint32_t)j + 3); g++) {
                boxes[nk] = gb[g];
                codes[nk++] = ~(int32_t)(((nn + g) << 3) | 0u);
            }
            boxes[nk] = below_box;
            codes[nk++] = below;
            const float inf = std::numeric_limits<float>::infinity();
            for (int i = 0; i < 4; i++) {
                bool used = i < nk;
                o[0 + i] = used ? boxes[i].lo[0] : inf;
                o[4 + i] = used ? boxes[i].hi[0] : -inf;
                o[8 + i] = used ? boxes[i].lo[1] : inf;
                o[12 + i] = used ? boxes[i].hi[1] : -inf;
                o[16 + i] = used ? boxes[i].lo[2] : inf;
                o[20 + i] = used ? boxes[i].hi[2] : -inf;
                int32_t c = used ? codes[i] : 0x7FFFFFFF;
                std::memcpy(&o[24 + i], &c, 4);
                o[28 + i] = 0.0f;
            }
            for (int i = 0; i + 1 < nk; i++) below_box.grow(boxes[i]);
            below = (int32_t)n4 + j;
        }
        root4 = below;
        depth4 += n_chain4;
        BD_CU(cudaMemcpy(nodes4 + 8 * (size_t)n4, chain4.data(), chain4.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    BD_CU(cudaEventRecord(e1, 0));
    BD_CU(cudaEventSynchronize(e1));
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    BD_CU(cudaGetLastError());
    DevScene &v = sd->view;
    v.fast_nodes = nodes2;
    v.fast_leaves = leaves;
    v.fast_root = root2;
    v.fast4_nodes = nodes4;
    v.fast4_root = root4;
    v.fast4_depth = depth4;
    info.ms = ms;
    info.nodes2 = (nn - 1) + n_chain2;
    info.nodes4 = n4 + n_chain4;
    info.depth2 = depth2;
    info.depth4 = depth4;
    info.giants = n_giants;
    info.leaves = n_slots;
    info.hidden = n - h.n_visible;
    return true;
}

// Copies the device-resident production tree back (tests: structural validation on the host).
bool download_fast_tree(const SceneDevice *sd, const DeviceTreeInfo &info, std::vector<float> &nodes2, std::vector<float> &nodes4,
                        std::vector<float> &leaves, std::string &err) {
    BD_CU(cudaSetDevice(sd->device));
    nodes2.resize(16 * (size_t)info.nodes2);
    nodes4.resize(32 * (size_t)info.nodes4);
    leaves.resize(12 * (size_t)info.leaves);
    BD_CU(cudaMemcpy(nodes2.data(), sd->view.fast_nodes, nodes2.size() * 4, cudaMemcpyDeviceToHost));
    BD_CU(cudaMemcpy(nodes4.data(), sd->view.fast4_nodes, nodes4.size() * 4, cudaMemcpyDeviceToHost));
    BD_CU(cudaMemcpy(leaves.data(), sd->view.fast_leaves, leaves.size() * 4, cudaMemcpyDeviceToHost));
    return true;
}

} // namespace paths
