// oracle.cpp -- CPU (f64) restatement of the render hot path of rynorris/paths.
//
// TEST INFRASTRUCTURE ONLY (see oracle.h).  Not part of the product; the product
// (paths_b200/) never links, loads or calls this file.
//
// Every function cites the reference lines (under /root/reference/) it follows.
// Arithmetic is f64 with the reference's operation order; build with
// -ffp-contract=off (rustc never contracts a*b+c into an FMA).
//
// Parity status: pinned only by the reference's own unit-test vectors (see
// tests/test_oracle_kat.py).  The reference draws every random number from an
// unseedable thread_rng (src/trace.rs:106, src/geom.rs:11,149, src/material.rs:310,
// src/scene.rs:201, src/sampling.rs:183), so no random stream can be matched; the
// oracle uses a counter-based Philox4x32-10 stream keyed by (seed, pixel, sample)
// -- the same stream definition the CUDA product uses, so that oracle and product
// follow the same paths up to f32/f64 rounding.
#include "oracle.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <vector>

namespace {

thread_local std::string g_err;

// ---------------------------------------------------------------- vector.rs
struct V3 {
    double x, y, z;
};
inline V3 v3(double x, double y, double z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }      // vector.rs:92-102
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }      // vector.rs:112-122
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }      // vector.rs:124-134
inline V3 operator*(V3 a, double s) { return v3(a.x * s, a.y * s, a.z * s); }        // vector.rs:136-147
inline V3 operator/(V3 a, double s) { return v3(a.x / s, a.y / s, a.z / s); }        // vector.rs:149-160
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }          // vector.rs:23-25
inline double magnitude(V3 a) { return dot(a, a); }                                  // vector.rs:27-29 (squared length!)
inline double vmax(V3 a) { return std::fmax(a.x, std::fmax(a.y, a.z)); }             // vector.rs:31-33 (NaN-ignoring f64::max)
inline double vmin(V3 a) { return std::fmin(a.x, std::fmin(a.y, a.z)); }             // vector.rs:35-37
inline V3 normed(V3 a) { return a / std::sqrt(magnitude(a)); }                       // vector.rs:39-41
inline V3 cross(V3 a, V3 b) {                                                        // vector.rs:43-49
    return v3(a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x);
}
inline bool is_nan(V3 a) { return std::isnan(a.x) || std::isnan(a.y) || std::isnan(a.z); }
inline V3 invert(V3 a) { return v3(1.0 / a.x, 1.0 / a.y, 1.0 / a.z); }               // vector.rs:63-65
inline V3 cw_max(V3 a, V3 b) { return v3(std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z)); }
inline V3 cw_min(V3 a, V3 b) { return v3(std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z)); }

// vector.rs:51-61
inline void form_basis(V3 n, V3 &i, V3 &j, V3 &k) {
    j = n;
    if (std::fabs(j.x) == 0.0)
        i = v3(1.0, 0.0, 0.0);
    else
        i = normed(cross(j, v3(0.0, 1.0, 0.0)));
    k = cross(i, j);
}
// geom.rs:26-28
inline V3 switch_basis(V3 v, V3 i, V3 j, V3 k) { return i * v.x + j * v.y + k * v.z; }

// ---------------------------------------------------------------- matrix.rs
struct M3 {
    double c[9];
    double get(int r, int col) const { return c[r * 3 + col]; }
};
inline M3 matmul(const M3 &a, const M3 &b) { // matrix.rs:78-92
    M3 out;
    for (int r = 0; r < 3; r++)
        for (int col = 0; col < 3; col++) {
            double v = 0.0;
            for (int k = 0; k < 3; k++) v += a.get(r, k) * b.get(k, col);
            out.c[r * 3 + col] = v;
        }
    return out;
}
inline V3 operator*(const M3 &m, V3 o) { // matrix.rs:94-104
    return v3(m.get(0, 0) * o.x + m.get(0, 1) * o.y + m.get(0, 2) * o.z,
              m.get(1, 0) * o.x + m.get(1, 1) * o.y + m.get(1, 2) * o.z,
              m.get(2, 0) * o.x + m.get(2, 1) * o.y + m.get(2, 2) * o.z);
}
inline M3 rotation_x(double a) { double s = std::sin(a), c = std::cos(a); return M3{{1.0, 0.0, 0.0, 0.0, c, -s, 0.0, s, c}}; }
inline M3 rotation_y(double a) { double s = std::sin(a), c = std::cos(a); return M3{{c, 0.0, s, 0.0, 1.0, 0.0, -s, 0.0, c}}; }
inline M3 rotation_z(double a) { double s = std::sin(a), c = std::cos(a); return M3{{c, -s, 0.0, s, c, 0.0, 0.0, 0.0, 1.0}}; }
// matrix.rs:30-35: rotation(yaw, pitch, roll) = Rx(pitch) * Ry(yaw) * Rz(roll)
inline M3 rotation(double yaw, double pitch, double roll) {
    return matmul(matmul(rotation_x(pitch), rotation_y(yaw)), rotation_z(roll));
}

// ---------------------------------------------------------------- colour.rs
struct Col {
    double r, g, b;
};
inline Col col(double r, double g, double b) { return Col{r, g, b}; }
inline Col operator+(Col a, Col b) { return col(a.r + b.r, a.g + b.g, a.b + b.b); }
inline Col operator*(Col a, Col b) { return col(a.r * b.r, a.g * b.g, a.b * b.b); }
inline Col operator*(Col a, double s) { return col(a.r * s, a.g * s, a.b * s); }
inline Col operator/(Col a, double s) { return col(a.r / s, a.g / s, a.b / s); }
inline double cmax(Col c) { // colour.rs:38-41
    double w = c.r > c.g ? c.r : c.g;
    return w > c.b ? w : c.b;
}
inline bool col_bad(Col c) { return c.r < 0.0 || c.g < 0.0 || c.b < 0.0; } // colour.rs:56-60 (panic there)
const Col BLACK{0.0, 0.0, 0.0};
const Col WHITE{1.0, 1.0, 1.0};

// colour.rs:62-74 (Rust `as u8` saturates, NaN -> 0)
inline uint8_t component_to_byte(double x) {
    x = std::pow(x, 0.45);
    if (x >= 1.0) return 255;
    if (x <= 0.0) return 0;
    double v = x * 256.0;
    if (std::isnan(v)) return 0;
    if (v >= 255.0) return 255;
    if (v <= 0.0) return 0;
    return (uint8_t)v;
}

// ---------------------------------------------------------------- geom.rs
struct Ray { // geom.rs:31-46
    V3 origin, direction, inv_direction;
};
inline Ray make_ray(V3 o, V3 d) { return Ray{o, d, invert(d)}; }

struct AABB { // geom.rs:72-83
    V3 min, max, center;
};
inline AABB make_aabb(V3 mn, V3 mx) { return AABB{mn, mx, (mn + mx) * 0.5}; }

struct Collision { // geom.rs:59-70
    double distance;
    V3 location, normal;
    bool has_meta;
    size_t face;
    double bx, by, bz;
};

struct Prim { // geom.rs:125-129
    int type; // 0 sphere, 1 triangle
    V3 center;
    double radius;
    size_t index;
    V3 v[3];
    V3 n;
};

struct Counters {
    uint64_t samples = 0, extend_rays = 0, shadow_rays = 0, box_tests = 0, tri_tests = 0, sphere_tests = 0, segments = 0,
             bad_samples = 0;
    void add(const Counters &o) {
        samples += o.samples; extend_rays += o.extend_rays; shadow_rays += o.shadow_rays; box_tests += o.box_tests;
        tri_tests += o.tri_tests; sphere_tests += o.sphere_tests; segments += o.segments; bad_samples += o.bad_samples;
    }
};

// geom.rs:208-235
inline bool sphere_intersect(const Prim &s, const Ray &ray, Collision &out) {
    V3 c = s.center;
    double r = s.radius;
    V3 o = ray.origin, l = ray.direction;
    double discriminant = (dot(l, o - c) * dot(l, o - c)) - dot(o - c, o - c) + (r * r);
    if (discriminant < 0.0) return false;
    double tmp = -dot(l, o - c);
    double sq = std::sqrt(discriminant);
    double d1 = tmp + sq;
    double d2 = tmp - sq;
    if (d1 < 0.0) return false;
    double distance = d2 > 0.0 ? d2 : d1;
    V3 location = o + (l * distance);
    V3 normal = normed(location - c);
    out = Collision{distance, location, normal, false, 0, 0.0, 0.0, 0.0};
    return true;
}
inline AABB sphere_aabb(const Prim &s) { // geom.rs:237-240
    V3 rv = v3(s.radius, s.radius, s.radius);
    return make_aabb(s.center - rv, s.center + rv);
}

// geom.rs:265-303
inline bool triangle_intersect(const Prim &tr, const Ray &ray, Collision &out) {
    V3 a = tr.v[0], b = tr.v[1], c = tr.v[2], n = tr.n;
    double cos_theta = dot(n, ray.direction);
    double d = dot(n, a);
    double t = (d - dot(n, ray.origin)) / cos_theta;
    if (std::isnan(t) || t < 0.0) return false;
    V3 p = ray.origin + (ray.direction * t);
    double area_abc = dot(n, cross(b - a, c - a));
    double area_pbc = dot(n, cross(b - p, c - p));
    double area_pca = dot(n, cross(c - p, a - p));
    double bx = area_pbc / area_abc;
    double by = area_pca / area_abc;
    double bz = 1.0 - bx - by;
    if (bx < 0.0 || by < 0.0 || bz < 0.0) return false;
    double back = cos_theta > 0.0 ? -1.0 : 1.0;
    out = Collision{t, p, n * back, true, tr.index, bx, by, bz};
    return true;
}
inline AABB triangle_aabb(const Prim &tr) { // geom.rs:305-320
    V3 v1 = tr.v[0], v2 = tr.v[1], v3_ = tr.v[2];
    double min_x = std::fmin(v1.x, std::fmin(v2.x, v3_.x));
    double min_y = std::fmin(v1.y, std::fmin(v2.y, v3_.y));
    double min_z = std::fmin(v1.z, std::fmin(v2.z, v3_.z));
    double max_x = std::fmax(v1.x, std::fmax(v2.x, v3_.x));
    double max_y = std::fmax(v1.y, std::fmax(v2.y, v3_.y));
    double max_z = std::fmax(v1.z, std::fmax(v2.z, v3_.z));
    return make_aabb(v3(min_x, min_y, min_z), v3(max_x, max_y, max_z));
}
inline bool prim_intersect(const Prim &p, const Ray &ray, Collision &out, Counters *ctr) { // geom.rs:184-189
    if (p.type == 0) {
        if (ctr) ctr->sphere_tests++;
        return sphere_intersect(p, ray, out);
    }
    if (ctr) ctr->tri_tests++;
    return triangle_intersect(p, ray, out);
}
inline AABB prim_aabb(const Prim &p) { return p.type == 0 ? sphere_aabb(p) : triangle_aabb(p); }

// geom.rs:251-261 (triangle), geom.rs:198-205 (sphere)
inline Prim prim_transform(const Prim &p, V3 translation, const M3 &rot, double scale) {
    Prim q = p;
    if (p.type == 0) {
        q.center = p.center + translation;
        q.radius = p.radius * scale;
    } else {
        for (int i = 0; i < 3; i++) q.v[i] = (rot * p.v[i]) * scale + translation;
        q.n = rot * p.n;
    }
    return q;
}

// ---------------------------------------------------------------- bvh.rs
// bvh.rs:8-21
inline bool ray_box_collide(const Ray &ray, const AABB &aabb, bool has_len, double len, double &tmin_out, Counters *ctr) {
    if (ctr) ctr->box_tests++;
    V3 t0s = (aabb.min - ray.origin) * ray.inv_direction;
    V3 t1s = (aabb.max - ray.origin) * ray.inv_direction;
    V3 tsmaller = cw_min(t0s, t1s);
    V3 tbigger = cw_max(t0s, t1s);
    double tmin = vmax(tsmaller);
    double tmax = vmin(tbigger);
    if (tmin < tmax && (!has_len || tmin < len)) {
        tmin_out = tmin;
        return true;
    }
    return false;
}

struct Node { // bvh.rs:23-70; Box<Node> pointers restated as indices into one arena
    bool leaf;
    int left, right;
    size_t obj;
    AABB aabb;
};

inline AABB combine_aabb(const AABB &a, const AABB &b) { // bvh.rs:352-366
    V3 mn = v3(std::fmin(a.min.x, b.min.x), std::fmin(a.min.y, b.min.y), std::fmin(a.min.z, b.min.z));
    V3 mx = v3(std::fmax(a.max.x, b.max.x), std::fmax(a.max.y, b.max.y), std::fmax(a.max.z, b.max.z));
    return make_aabb(mn, mx);
}
inline double surface_area(const AABB &a) { // bvh.rs:368-373
    double w = a.max.x - a.min.x, h = a.max.y - a.min.y, d = a.max.z - a.min.z;
    return 2.0 * (w * h + h * d + d * w);
}

const size_t DELTA = 10;     // bvh.rs:149
const double EPSILON = 0.01; // bvh.rs:153
inline size_t ccrf(size_t x) { // bvh.rs:154-158
    double xf = (double)x;
    double c = std::pow((double)DELTA, 0.5 - EPSILON) / 2.0;
    double v = std::ceil(c * std::pow(xf, 0.5 - EPSILON));
    return (size_t)v;
}

inline uint64_t morton_code(uint16_t num_bits, uint16_t x, uint16_t y, uint16_t z) { // bvh.rs:373-384
    uint64_t mc = 0;
    for (uint16_t ix = 0; ix < num_bits; ix++) {
        mc |= ((uint64_t)(z & 1)) << (64 - (num_bits * 3) + (ix * 3));
        mc |= ((uint64_t)(y & 1)) << (64 - (num_bits * 3) + (ix * 3) + 1);
        mc |= ((uint64_t)(x & 1)) << (64 - (num_bits * 3) + (ix * 3) + 2);
        x >>= 1; y >>= 1; z >>= 1;
    }
    return mc;
}
inline bool get_bit(uint64_t mc, uint16_t bit) { return ((mc >> (63 - bit)) & 1) == 1; } // bvh.rs:386-388

inline uint16_t sat_u16(double v) { // Rust `as u16`: saturating, NaN -> 0
    if (std::isnan(v)) return 0;
    if (v <= 0.0) return 0;
    if (v >= 65535.0) return 65535;
    return (uint16_t)v;
}
inline uint32_t sat_u32(double v) { // Rust `as u32`
    if (std::isnan(v)) return 0;
    if (v <= 0.0) return 0;
    if (v >= 4294967295.0) return 4294967295u;
    return (uint32_t)v;
}

struct BVH {
    std::vector<Node> nodes; // arena, capacity 2N-1, filled through `next`
    std::atomic<int> next{0};
    std::vector<Prim> prims; // leaf.primitive, indexed by leaf.obj
    int root = -1;

    int alloc(const Node &n) {
        int ix = next.fetch_add(1);
        nodes[ix] = n;
        return ix;
    }
    const AABB &aabb(int n) const { return nodes[n].aabb; }
    double cost(int a, int b) const { return surface_area(combine_aabb(aabb(a), aabb(b))); } // bvh.rs:345-350

    size_t find_best_match(const std::vector<int> &clusters, size_t ix) const { // bvh.rs:327-343
        double lowest = std::numeric_limits<double>::max();
        size_t best_jx = 0;
        for (size_t jx = 0; jx < clusters.size(); jx++) {
            if (ix == jx) continue;
            double c = cost(clusters[ix], clusters[jx]);
            if (c < lowest) {
                lowest = c;
                best_jx = jx;
            }
        }
        return best_jx;
    }

    // bvh.rs:269-325 (including the index fix-up loop, verbatim in behaviour)
    std::vector<int> combine_clusters(std::vector<int> clusters, size_t n) {
        std::vector<size_t> closest;
        closest.reserve(clusters.size());
        for (size_t ix = 0; ix < clusters.size(); ix++) closest.push_back(find_best_match(clusters, ix));
        while (clusters.size() > n) {
            double best = std::numeric_limits<double>::max();
            size_t left = 0, right = 0;
            for (size_t ix = 0; ix < clusters.size(); ix++) {
                double c = cost(clusters[ix], clusters[closest[ix]]);
                if (c < best) {
                    best = c;
                    left = ix;
                    right = closest[ix];
                }
            }
            if (right < left) std::swap(right, left);
            int lc = clusters[right];
            clusters.erase(clusters.begin() + right);
            int rc = clusters[left];
            clusters.erase(clusters.begin() + left);
            closest.erase(closest.begin() + right);
            closest.erase(closest.begin() + left);
            Node comb; // ClusterNode::new(Box::new(lc), Box::new(rc)), bvh.rs:55-70
            comb.leaf = false;
            comb.left = lc;
            comb.right = rc;
            comb.obj = 0;
            comb.aabb = combine_aabb(aabb(lc), aabb(rc));
            clusters.push_back(alloc(comb));
            closest.push_back(find_best_match(clusters, clusters.size() - 1));
            for (size_t ix = 0; ix < clusters.size(); ix++) {
                if (closest[ix] == left || closest[ix] == right)
                    closest[ix] = find_best_match(clusters, ix);
                else if (closest[ix] >= right)
                    closest[ix] -= 2;
                else if (closest[ix] >= left)
                    closest[ix] -= 1;
            }
        }
        return clusters;
    }

    typedef std::vector<std::pair<int, uint64_t>> MCVec;

    // bvh.rs:240-267
    static void make_partition(MCVec &clusters, uint16_t depth, MCVec &lhs, MCVec &rhs) {
        if (clusters.empty()) return;
        if (get_bit(clusters.front().second, depth)) { rhs.swap(clusters); return; }
        if (!get_bit(clusters.back().second, depth)) { lhs.swap(clusters); return; }
        size_t max_0 = 0, min_1 = clusters.size() - 1;
        while (min_1 - max_0 > 1) {
            size_t mid = (min_1 + max_0) / 2;
            if (get_bit(clusters[mid].second, depth)) min_1 = mid; else max_0 = mid;
        }
        rhs.assign(clusters.begin() + min_1, clusters.end());
        clusters.resize(min_1);
        lhs.swap(clusters);
    }

    // bvh.rs:208-238
    std::vector<int> build_tree(MCVec clusters, uint16_t max_depth, uint16_t depth) {
        size_t num_clusters = clusters.size();
        if (num_clusters < DELTA) {
            std::vector<int> plain;
            for (auto &c : clusters) plain.push_back(c.first);
            return combine_clusters(plain, ccrf(DELTA));
        }
        MCVec lhs, rhs;
        if (depth < max_depth) {
            make_partition(clusters, depth, lhs, rhs);
        } else {
            size_t mid = clusters.size() / 2;
            rhs.assign(clusters.begin() + mid, clusters.end());
            clusters.resize(mid);
            lhs.swap(clusters);
        }
        std::vector<int> new_clusters;
        if (depth < 2) { // bvh.rs:223-231 forks two threads; result order is left then right either way
            std::vector<int> l, r;
            std::thread tl([&] { l = build_tree(std::move(lhs), max_depth, depth + 1); });
            std::thread trr([&] { r = build_tree(std::move(rhs), max_depth, depth + 1); });
            tl.join();
            trr.join();
            new_clusters = l;
            new_clusters.insert(new_clusters.end(), r.begin(), r.end());
        } else {
            new_clusters = build_tree(std::move(lhs), max_depth, depth + 1);
            std::vector<int> r = build_tree(std::move(rhs), max_depth, depth + 1);
            new_clusters.insert(new_clusters.end(), r.begin(), r.end());
        }
        return combine_clusters(new_clusters, ccrf(num_clusters));
    }

    // bvh.rs:160-206
    bool construct(const std::vector<Prim> &items) {
        prims = items;
        size_t n = items.size();
        if (n == 0) { g_err = "Must have at least one cluster (empty scene)"; return false; }
        nodes.assign(2 * n, Node());
        next = 0;
        std::vector<int> leaves;
        for (size_t ix = 0; ix < n; ix++) {
            Node leaf;
            leaf.leaf = true;
            leaf.left = leaf.right = -1;
            leaf.obj = ix;
            leaf.aabb = prim_aabb(items[ix]);
            leaves.push_back(alloc(leaf));
        }
        uint16_t num_bits = sat_u16(std::ceil(std::log((double)n) / std::log(4.0))); // f64::log(4.0) = ln/ln
        if (num_bits > 16) { g_err = "Too many objects to construct BVH"; return false; }
        double cap = (double)(1 << num_bits);
        double mx = std::numeric_limits<double>::quiet_NaN();
        for (int l : leaves) mx = std::fmax(mx, vmax(aabb(l).center));
        double scale = cap / mx;
        MCVec with_mc;
        for (int l : leaves) {
            V3 c = aabb(l).center;
            uint64_t mc = morton_code(num_bits, sat_u16(c.x * scale), sat_u16(c.y * scale), sat_u16(c.z * scale));
            with_mc.push_back(std::make_pair(l, mc));
        }
        // bvh.rs:191 uses sort_unstable_by_key whose tie order is unspecified; the oracle keeps input order on ties.
        std::stable_sort(with_mc.begin(), with_mc.end(),
                         [](const std::pair<int, uint64_t> &a, const std::pair<int, uint64_t> &b) { return a.second < b.second; });
        std::vector<int> clusters = build_tree(std::move(with_mc), num_bits, 0);
        std::vector<int> fin = combine_clusters(clusters, 1);
        root = fin.back();
        return true;
    }

    // bvh.rs:78-140. Returns leaf.obj or -1.
    long find_intersection(const Ray &ray, Collision &best, Counters *ctr) const {
        const int STACK = 100;
        int stack[STACK];
        int sp = 0;
        double dummy;
        if (!ray_box_collide(ray, aabb(root), false, 0.0, dummy, ctr)) return -1;
        int node = root;
        long best_obj = -1;
        for (;;) {
            const Node &nd = nodes[node];
            if (nd.leaf) {
                Collision colr;
                if (prim_intersect(prims[nd.obj], ray, colr, ctr)) {
                    if (best_obj >= 0) {
                        if (colr.distance < best.distance) { best = colr; best_obj = (long)nd.obj; }
                    } else {
                        best = colr;
                        best_obj = (long)nd.obj;
                    }
                }
                if (sp == 0) break;
                node = stack[--sp];
            } else {
                double ld = 0.0, rd = 0.0;
                bool has = best_obj >= 0;
                bool lh = ray_box_collide(ray, aabb(nd.left), has, has ? best.distance : 0.0, ld, ctr);
                bool rh = ray_box_collide(ray, aabb(nd.right), has, has ? best.distance : 0.0, rd, ctr);
                if (lh && rh) {
                    if (sp >= STACK) { return best_obj; } // reference would panic (index OOB)
                    if (ld < rd) { stack[sp++] = nd.right; node = nd.left; }
                    else { stack[sp++] = nd.left; node = nd.right; }
                } else if (lh) node = nd.left;
                else if (rh) node = nd.right;
                else {
                    if (sp == 0) break;
                    node = stack[--sp];
                }
            }
        }
        return best_obj;
    }

    uint32_t depth_of(int n) const {
        if (nodes[n].leaf) return 1;
        return 1 + std::max(depth_of(nodes[n].left), depth_of(nodes[n].right));
    }
};

// ---------------------------------------------------------------- model.rs
struct Model {
    std::vector<V3> vertices;
    std::vector<uint32_t> faces; // 3 per face
    std::vector<V3> face_normals;
    std::vector<V3> vertex_normals; // empty until computed
    bool has_vn = false;
    std::vector<Col> vertex_colours;
    bool has_vc = false;

    static V3 face_normal(V3 v1, V3 v2, V3 v3_) { // model.rs:237-249
        V3 side_1 = v2 - v1, side_2 = v3_ - v1, side_3 = v3_ - v2;
        V3 n = normed(cross(side_1, side_2));
        if (std::isnan(n.x) || std::isnan(n.y) || std::isnan(n.z)) n = normed(cross(side_1, side_3));
        return n;
    }
    void compute_face_normals() { // model.rs:226-235
        size_t nf = faces.size() / 3;
        face_normals.resize(nf);
        for (size_t f = 0; f < nf; f++)
            face_normals[f] = face_normal(vertices[faces[3 * f]], vertices[faces[3 * f + 1]], vertices[faces[3 * f + 2]]);
    }
    void compute_vertex_normals() { // model.rs:194-224 (un-normalised mean of adjacent face normals)
        if (has_vn) return;
        std::vector<V3> sums(vertices.size(), v3(0.0, 0.0, 0.0));
        std::vector<size_t> counts(vertices.size(), 0);
        size_t nf = faces.size() / 3;
        for (size_t f = 0; f < nf; f++) {
            V3 n = face_normals[f];
            if (is_nan(n)) continue;
            for (int k = 0; k < 3; k++) {
                uint32_t v = faces[3 * f + k];
                sums[v] = sums[v] + n;
                counts[v] += 1;
            }
        }
        vertex_normals.resize(vertices.size());
        for (size_t i = 0; i < vertices.size(); i++) vertex_normals[i] = sums[i] / (double)counts[i];
        has_vn = true;
    }
    V3 smooth_normal(size_t face, double bx, double by, double bz) const { // model.rs:142-156
        V3 an = vertex_normals[faces[3 * face]], bn = vertex_normals[faces[3 * face + 1]], cn = vertex_normals[faces[3 * face + 2]];
        return an * bx + bn * by + cn * bz;
    }
    Col smooth_colour(size_t face, double bx, double by, double bz) const { // model.rs:158-172
        Col ac = vertex_colours[faces[3 * face]], bc = vertex_colours[faces[3 * face + 1]], cc = vertex_colours[faces[3 * face + 2]];
        return ac * bx + bc * by + cc * bz;
    }
    std::vector<Prim> resolve_primitives() const { // model.rs:174-192
        std::vector<Prim> out;
        size_t nf = faces.size() / 3;
        for (size_t f = 0; f < nf; f++) {
            V3 n = face_normals[f];
            if (is_nan(n)) continue;
            Prim p;
            p.type = 1;
            p.center = v3(0, 0, 0);
            p.radius = 0;
            p.index = f;
            p.v[0] = vertices[faces[3 * f]];
            p.v[1] = vertices[faces[3 * f + 1]];
            p.v[2] = vertices[faces[3 * f + 2]];
            p.n = n;
            out.push_back(p);
        }
        return out;
    }
};

// ---------------------------------------------------------------- material.rs
const double PI = 3.14159265358979323846; // std::f64::consts::PI

struct BasicMat {
    int type;
    bool vertex_albedo;
    Col albedo;
    double r0, metalness, roughness;
};
struct Mat {
    BasicMat m;
    bool fresnel;
    double fresnel_r0;
    BasicMat diffuse, specular;
};

inline BasicMat to_basic(const orc_basic_material &d) {
    return BasicMat{d.type, d.albedo_mode == 1, col(d.albedo[0], d.albedo[1], d.albedo[2]), d.reflectance, d.metalness, d.roughness};
}
inline Mat to_mat(const orc_material_desc &d) {
    Mat m;
    m.m = to_basic(d.m);
    m.fresnel = d.m.type == ORC_MAT_FRESNEL;
    m.diffuse = to_basic(d.diffuse);
    m.specular = to_basic(d.specular);
    // material.rs:380-387
    double n1 = 1.0, n2 = d.refractive_index;
    m.fresnel_r0 = std::pow((n1 - n2) / (n1 + n2), 2.0);
    return m;
}

inline V3 reflect(V3 vector, V3 normal) { // material.rs:249-251
    return normed((normal * dot(normal, vector) * 2.0) - vector);
}
// geom.rs:10-24 with the two uniforms supplied (u first, then v)
inline V3 cosine_sample_hemisphere(double u, double v) {
    double r = std::sqrt(u);
    double theta = 2.0 * PI * v;
    double x = r * std::cos(theta);
    double y = 1.0 - u;
    double z = r * std::sin(theta);
    return v3(x, y, z);
}
inline V3 lambert_sample_pdf(V3 normal, double u, double v) { // material.rs:224-231
    V3 rd = cosine_sample_hemisphere(u, v);
    V3 i, j, k;
    form_basis(normal, i, j, k);
    return normed(switch_basis(rd, i, j, k));
}
inline double lambert_weight_pdf(V3 vec_in, V3 normal) { return dot(normal, vec_in * -1.0) / PI; }          // material.rs:220-222
inline Col lambert_brdf(Col albedo, V3 vec_in, V3 normal) { return albedo * dot(normal, vec_in * -1.0) / PI; } // material.rs:237-239
inline double schlick(double r0, V3 vec_out, V3 normal) { // material.rs:303-305
    double cos_theta = dot(vec_out, normal);
    return r0 + (1.0 - r0) * std::pow(1.0 - cos_theta, 5.0);
}
inline double ct_ndf(double m, V3 n, V3 h) { // material.rs:436-447 (Beckmann)
    double alpha = std::acos(dot(h, n));
    double cos_alpha = std::cos(alpha);
    double tan_alpha = std::tan(alpha);
    double e = std::exp(-1.0 * (tan_alpha * tan_alpha) / (m * m));
    double d0 = e / (PI * m * m * std::pow(cos_alpha, 4.0));
    return std::fmax(0.0, d0 * dot(n, h));
}
inline Col basic_brdf(const BasicMat &b, V3 vec_out, V3 vec_in, V3 normal) {
    switch (b.type) {
    case ORC_MAT_LAMBERTIAN: return lambert_brdf(b.albedo, vec_in, normal);
    case ORC_MAT_MIRROR: return BLACK; // material.rs:268-271
    case ORC_MAT_GLOSS: { // material.rs:360-370
        double r = schlick(b.r0, vec_out, normal);
        Col diffuse = lambert_brdf(b.albedo, vec_in, normal) * (1.0 - b.metalness);
        Col specular = BLACK;
        return diffuse * (1.0 - r) + specular * r;
    }
    case ORC_MAT_COOK_TORRANCE: { // material.rs:503-523
        V3 h = normed(vec_out - vec_in);
        double d = ct_ndf(b.roughness, normal, h);
        double ndl = dot(normal, vec_in * -1.0);
        double vdh = dot(vec_out, h);
        double ndh = dot(normal, h);
        double ndv = dot(normal, vec_out);
        double g = std::fmax(0.0, std::fmin(1.0, std::fmin((2.0 * ndh * ndv) / vdh, (2.0 * ndh * ndl) / vdh)));
        return b.albedo * (d * g) / (4.0 * ndv * ndl);
    }
    }
    return BLACK;
}
inline Col mat_brdf(const Mat &m, V3 vec_out, V3 vec_in, V3 normal) { // material.rs:120-128, 414-421
    if (m.fresnel) {
        double r = schlick(m.fresnel_r0, vec_out, normal);
        return basic_brdf(m.diffuse, vec_out, vec_in, normal) * (1.0 - r) + basic_brdf(m.specular, vec_out, vec_in, normal) * r;
    }
    return basic_brdf(m.m, vec_out, vec_in, normal);
}
// material.rs:81-88 -- only Lambertian / Mirror / Gloss can be sampled; the rest panics in the reference.
// u[0] = gloss choice, u[1], u[2] = hemisphere (u, v).  Returns false if the material cannot be sampled.
inline bool mat_sample(const Mat &m, V3 vec_out, V3 normal, const double u[3], V3 &direction, double &pdf, Col &brdf,
                       bool &is_specular) {
    if (m.fresnel) return false;
    const BasicMat &b = m.m;
    if (b.type == ORC_MAT_LAMBERTIAN) { // material.rs:211-216
        direction = lambert_sample_pdf(normal, u[1], u[2]);
        pdf = lambert_weight_pdf(direction * -1.0, normal);
        brdf = lambert_brdf(b.albedo, direction * -1.0, normal);
        is_specular = false;
        return true;
    }
    if (b.type == ORC_MAT_MIRROR) { // material.rs:253-255
        direction = reflect(vec_out, normal);
        pdf = 1.0;
        brdf = WHITE;
        is_specular = true;
        return true;
    }
    if (b.type == ORC_MAT_GLOSS) { // material.rs:302-325
        double r = schlick(b.r0, vec_out, normal);
        double specular_chance = b.r0 > 0.5 ? r : 0.5;
        is_specular = u[0] <= specular_chance;
        if (is_specular) {
            direction = reflect(vec_out, normal);
            double p = 1.0;
            Col br = b.albedo * b.metalness + WHITE * (1.0 - b.metalness);
            pdf = p * specular_chance;
            brdf = br * r;
        } else {
            direction = lambert_sample_pdf(normal, u[1], u[2]);
            V3 vec_in = direction * -1.0;
            double p = lambert_weight_pdf(vec_in, normal);
            Col br = lambert_brdf(b.albedo, vec_in, normal) * (1.0 - b.metalness);
            pdf = p * (1.0 - specular_chance);
            brdf = br * (1.0 - r);
        }
        return true;
    }
    return false;
}

// ---------------------------------------------------------------- RNG (oracle's own; shared definition with the product)
inline void philox4x32_10(const uint32_t c_in[4], const uint32_t k_in[2], uint32_t out[4]) {
    uint32_t c0 = c_in[0], c1 = c_in[1], c2 = c_in[2], c3 = c_in[3];
    uint32_t k0 = k_in[0], k1 = k_in[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
const uint32_t TAG_PATH = 0x50415448u; // 'PATH'
const uint32_t TAG_CMJ = 0x434D4A20u;  // 'CMJ '
inline double u01(uint32_t x) { return (double)(x >> 8) * (1.0 / 16777216.0); } // 24-bit uniform in [0,1)

struct PathRng { // dims per bounce: block 0 = {light choice, light u, light v, gloss choice}; block 1 = {hemi u, hemi v, rr, spare}
    uint32_t key[2], pixel, sample;
    void block(uint32_t bounce, uint32_t k, uint32_t out[4]) const {
        uint32_t ctr[4] = {pixel, sample, bounce * 2 + k, TAG_PATH};
        philox4x32_10(ctr, key, out);
    }
};

// ---------------------------------------------------------------- sampling.rs (CMJ)
inline uint32_t cmj_permute(uint32_t i, uint32_t l, uint32_t p) { // sampling.rs:187-210 (note `while i > l`)
    uint32_t w = l - 1;
    w |= w >> 1; w |= w >> 2; w |= w >> 4; w |= w >> 8; w |= w >> 16;
    while (i > l) {
        i ^= p;             i *= 0xe170893du;
        i ^= p >> 16;
        i ^= (i & w) >> 4;
        i ^= p >> 8;        i *= 0x0929eb3fu;
        i ^= p >> 23;
        i ^= (i & w) >> 1;  i *= (1u | p >> 27);
                            i *= 0x6935fa69u;
        i ^= (i & w) >> 11; i *= 0x74dcb303u;
        i ^= (i & w) >> 2;  i *= 0x9e501cc3u;
        i ^= (i & w) >> 2;  i *= 0xc860a3dfu;
        i &= w;
        i ^= i >> 5;
    }
    return (i + p) % l; // wrapping add in release builds
}
inline double cmj_rand_float(uint32_t i, uint32_t p) { // sampling.rs:212-221
    i ^= p;
    i ^= i >> 17;
    i ^= i >> 10; i *= 0xb36534e5u;
    i ^= i >> 12;
    i ^= i >> 21; i *= 0x93fc4795u;
    i ^= 0xdf6e307fu;
    i ^= i >> 17; i *= (1u | p >> 18);
    return (double)i * (1.0 / 4294967808.0);
}
inline void cmj(uint32_t s, uint32_t m, uint32_t n, uint32_t p, double &x, double &y) { // sampling.rs:226-235
    uint32_t ps = cmj_permute(s, m * n, p * 0xa73bd290u);
    double sx = (double)cmj_permute(ps % m, m, p * 0xa511e9b3u);
    double sy = (double)cmj_permute(ps / m, n, p * 0x63d83595u);
    double jx = cmj_rand_float(s, p * 0xa399d265u);
    double jy = cmj_rand_float(s, p * 0x711ad6a5u);
    x = (((double)(s % m)) + (sy + jx) / ((double)n)) / ((double)m);
    y = (((double)(s / m)) + (sx + jy) / ((double)m)) / ((double)n);
}
inline void cmj_disk(uint32_t s, uint32_t m, uint32_t n, uint32_t p, double &dx, double &dy) { // sampling.rs:250-264
    double x, y;
    cmj(s, m, n, p, x, y);
    double theta = 2.0 * PI * x;
    double r = std::sqrt(y);
    dx = r * std::cos(theta);
    dy = r * std::sin(theta);
}
// renderer.rs:166-192 + worker.rs:58-87: the first sample of a pixel comes from a (1,1) pattern, every later
// one from successive (5,5) patterns.  out = {m, n, s, pattern number}.
inline void sample_schedule(uint32_t S, uint32_t &m, uint32_t &n, uint32_t &s, uint32_t &pattern) {
    if (S == 0) { m = 1; n = 1; s = 0; pattern = 0; }
    else { m = 5; n = 5; s = (S - 1) % 25; pattern = 1 + (S - 1) / 25; }
}

// ---------------------------------------------------------------- camera.rs
struct Camera {
    V3 location;
    double focal_length, distance_from_lens, aperture;
    M3 rot;
    double sensor_width, sensor_height;
    uint32_t width, height;
};
inline Camera to_camera(const orc_camera_desc &d) { // serde.rs:172-188
    Camera c;
    c.width = d.image_width;
    c.height = d.image_height;
    c.location = v3(d.location[0], d.location[1], d.location[2]);
    c.rot = rotation(d.yaw, d.pitch, d.roll);
    c.sensor_width = d.sensor_width;
    c.sensor_height = d.sensor_height;
    c.focal_length = d.focal_length;
    c.aperture = d.aperture;
    c.distance_from_lens = (d.focal_length * d.focus_distance) / (d.focus_distance - d.focal_length);
    return c;
}
// camera.rs:47-94
inline void get_ray_for_pixel(const Camera &cam, uint32_t x, uint32_t y, double sqx, double sqy, double dkx, double dky, Ray &ray,
                              double &weight) {
    x = cam.width - x - 1;
    y = cam.height - y - 1;
    double f = cam.focal_length, v = cam.distance_from_lens;
    double p = (f * v) / (v - f);
    double x_scale = cam.sensor_width / ((double)cam.width);
    double y_scale = cam.sensor_height / ((double)cam.height);
    double image_x = ((double)x) - ((double)cam.width) / 2.0 + sqx;
    double image_y = ((double)cam.height) / 2.0 - ((double)y) - sqy;
    V3 k = v3(image_x * x_scale, image_y * y_scale, -cam.distance_from_lens);
    double aperture_radius = cam.focal_length / cam.aperture; // camera.rs:41-45
    V3 l = v3(dkx * aperture_radius, dky * aperture_radius, 0.0);
    V3 dir = ((k * (p / v)) + l) * -1.0;
    V3 norm_dir = normed(dir);
    V3 origin = cam.rot * l + cam.location;
    V3 direction = cam.rot * norm_dir;
    weight = norm_dir.z;
    ray = make_ray(origin, direction);
}

// ---------------------------------------------------------------- scene.rs
struct Object {
    int shape; // 0 sphere, 1 mesh
    int model;
    bool smooth;
    M3 rot;
    Mat material;
};
struct Light {
    int geometry; // 0 point, 1 area sphere
    V3 v;
    double radius;
    Col colour;
    double intensity;
};
struct Entity {
    bool is_light;
    uint32_t id;
};

} // namespace

struct orc_scene {
    std::vector<Model> models;
    std::vector<Object> objects;
    std::vector<Light> lights;
    std::vector<Entity> items; // BVH<EntityID>::items, indexed by leaf.obj
    BVH bvh;
    int sky_type;
    Col sky_c0, sky_c1;
    uint32_t sky_w, sky_h;
    std::vector<float> sky_texels;
};

namespace {

// scene.rs:88-113
inline Col ambient_light(const orc_scene &s, V3 direction) {
    if (s.sky_type == 0) return s.sky_c0;
    if (s.sky_type == 1) {
        double cos_theta = dot(direction, v3(0.0, 1.0, 0.0));
        return s.sky_c0 * cos_theta + s.sky_c1 * (1.0 - cos_theta);
    }
    double lat = std::acos(direction.y);
    double lon = std::atan2(direction.z, direction.x);
    double w = (double)s.sky_w, h = (double)s.sky_h;
    double x = w / 2.0 * (lon / PI) + w / 2.0;
    double y = h * (1.0 - lat / PI);
    uint32_t x_pix = std::min(s.sky_w - 1, sat_u32(x));
    uint32_t y_pix = std::min(s.sky_h - 1, sat_u32(y));
    size_t pix = (size_t)y_pix * s.sky_w + x_pix;
    return col((double)s.sky_texels[3 * pix], (double)s.sky_texels[3 * pix + 1], (double)s.sky_texels[3 * pix + 2]);
}

// geom.rs:146-173 with the two uniforms supplied (u then v)
inline void sphere_light_sample(V3 center, double radius, V3 from, double u, double v, V3 &out_dir, double &inv_pdf) {
    double theta = 2.0 * PI * u;
    double phi = std::acos(2.0 * v - 1.0);
    V3 n = v3(std::sin(phi) * std::cos(theta), std::sin(phi) * std::sin(theta), std::cos(phi));
    V3 point = center + n * radius;
    V3 out_vec = from - point;
    out_dir = normed(out_vec);
    double distance_sq = magnitude(out_vec);
    double area = 4.0 * PI * radius * radius;
    double ip = area * dot(n, out_dir) / distance_sq;
    inv_pdf = std::fmax(0.0, ip);
}
inline void light_sample(const Light &l, V3 from, double u, double v, V3 &dir, double &inv_pdf) { // scene.rs:54-59
    if (l.geometry == 0) { dir = l.v; inv_pdf = 1.0; }
    else sphere_light_sample(l.v, l.radius, from, u, v, dir, inv_pdf);
}

// scene.rs:172-197.  Returns item index or -1; normal replaced by the (un-normalised, un-flipped) smooth normal.
inline long scene_find_intersection(const orc_scene &s, const Ray &ray, Collision &c, Counters *ctr) {
    long obj = s.bvh.find_intersection(ray, c, ctr);
    if (obj < 0) return -1;
    const Entity &e = s.items[obj];
    if (!e.is_light) {
        const Object &o = s.objects[e.id];
        if (o.shape == 1 && c.has_meta && o.smooth) {
            V3 sn = s.models[o.model].smooth_normal(c.face, c.bx, c.by, c.bz);
            c.normal = o.rot * sn;
        }
    }
    return obj;
}

// trace.rs:7-121
inline Col trace_ray(const orc_scene &s, Ray ray, const PathRng &rng, Counters &ctr) {
    Col throughput = WHITE, colour = BLACK;
    int loops = 0;
    bool last_bounce_specular = true;
    for (;;) {
        if (loops > 10) break;
        Collision c;
        ctr.extend_rays++;
        long item = scene_find_intersection(s, ray, c, &ctr);
        ctr.segments++;
        if (item < 0) {
            colour = colour + throughput * ambient_light(s, ray.direction * -1.0);
            break;
        }
        double cos_in = dot(ray.direction, c.normal * -1.0);
        if (cos_in <= 0.0) break;
        const Entity &e = s.items[item];
        if (e.is_light) { // trace.rs:31-41
            const Light &l = s.lights[e.id];
            if (last_bounce_specular) {
                colour = colour + throughput * l.colour * l.intensity;
                if (col_bad(colour)) ctr.bad_samples++;
            }
            break;
        }
        const Object &o = s.objects[e.id];
        Mat material = o.material; // trace.rs:44-50 -> material.rs:39-49,183-195
        if (o.shape == 1) {
            if (!material.fresnel && (material.m.type == ORC_MAT_LAMBERTIAN || material.m.type == ORC_MAT_GLOSS) &&
                material.m.vertex_albedo) {
                material.m.albedo = s.models[o.model].smooth_colour(c.face, c.bx, c.by, c.bz);
                material.m.vertex_albedo = false;
            }
        }
        uint32_t b0[4], b1[4];
        rng.block((uint32_t)loops, 0, b0);
        rng.block((uint32_t)loops, 1, b1);

        // Next event estimation, trace.rs:53-82
        Col direct = BLACK;
        if (!s.lights.empty()) {
            size_t nl = s.lights.size();
            size_t id = std::min(nl - 1, (size_t)(u01(b0[0]) * (double)nl)); // scene.rs:199-206 gen_range(0, n)
            const Light &light = s.lights[id];
            V3 in_dir;
            double inv_pdf;
            light_sample(light, c.location, u01(b0[1]), u01(b0[2]), in_dir, inv_pdf);
            Ray shadow_ray = make_ray(c.location + c.normal * 0.0001, in_dir * -1.0);
            Collision sc;
            ctr.shadow_rays++;
            long sitem = scene_find_intersection(s, shadow_ray, sc, &ctr);
            bool occluded = false;
            if (sitem >= 0) {
                const Entity &se = s.items[sitem];
                occluded = !(se.is_light && se.id == (uint32_t)id);
            }
            double cos_theta = std::fmax(0.0, dot(c.normal, shadow_ray.direction));
            if (!(occluded || cos_theta <= 0.0)) {
                Col base = light.colour * light.intensity;
                Col brdf = mat_brdf(material, ray.direction * -1.0, shadow_ray.direction * -1.0, c.normal);
                direct = base * brdf * inv_pdf;
            }
        }
        if (col_bad(direct)) ctr.bad_samples++;
        colour = colour + direct * throughput;
        if (col_bad(colour)) ctr.bad_samples++;

        V3 direction;
        double pdf;
        Col brdf;
        bool is_specular;
        double us[3] = {u01(b0[3]), u01(b1[0]), u01(b1[1])};
        if (!mat_sample(material, ray.direction * -1.0, c.normal, us, direction, pdf, brdf, is_specular)) {
            ctr.bad_samples++; // reference panics ("Not implemented", material.rs:86)
            break;
        }
        last_bounce_specular = is_specular;
        Ray new_ray = make_ray(c.location + c.normal * 0.0001, direction);
        Col attenuation = brdf / pdf;
        throughput = throughput * attenuation;
        if (cmax(throughput) <= 0.0) break;
        // emittance is always BLACK for every material serde.rs can build (serde.rs:278-291); trace.rs:100-101 adds 0.
        if (loops >= 2) { // trace.rs:104-111
            double survival_chance = cmax(throughput);
            if (u01(b1[2]) > survival_chance) break;
            throughput = throughput / survival_chance;
        }
        ray = new_ray;
        loops += 1;
    }
    return colour;
}

inline void pixel_sample_ray(const Camera &cam, const uint32_t key[2], uint32_t x, uint32_t y, uint32_t S, Ray &ray, double &weight) {
    uint32_t m, n, s, pattern;
    sample_schedule(S, m, n, s, pattern);
    uint32_t pixel = x + y * cam.width;
    uint32_t ctr[4] = {pixel, pattern, 0, TAG_CMJ};
    uint32_t seeds[4];
    philox4x32_10(ctr, key, seeds); // worker.rs:69-70: one seed for the sensor pattern, one for the lens pattern
    double sqx, sqy, dkx, dky;
    cmj(s, m, n, seeds[0], sqx, sqy);
    cmj_disk(s, m, n, seeds[1], dkx, dky);
    get_ray_for_pixel(cam, x, y, sqx, sqy, dkx, dky, ray, weight);
}

inline Col trace_pixel(const orc_scene &s, const Camera &cam, const uint32_t key[2], uint32_t x, uint32_t y, uint32_t S, Counters &ctr) {
    Ray ray;
    double weight;
    pixel_sample_ray(cam, key, x, y, S, ray, weight);
    PathRng rng{{key[0], key[1]}, x + y * cam.width, S};
    ctr.samples++;
    return trace_ray(s, ray, rng, ctr) * weight; // worker.rs:76
}

inline bool owns_pixel(int mode, uint32_t rank, uint32_t world, uint32_t x, uint32_t y, uint32_t W) {
    if (mode == 0 || world <= 1) return true;
    uint32_t tiles_x = (W + 31) / 32; // the product's shard tiles (PATHS_SHARD_TILE): test infrastructure mirrors them
    uint32_t tile = (y / 32) * tiles_x + (x / 32);
    return tile % world == rank;
}

inline void export_counters(const Counters &c, orc_counters *o) {
    if (!o) return;
    o->samples += c.samples; o->extend_rays += c.extend_rays; o->shadow_rays += c.shadow_rays; o->box_tests += c.box_tests;
    o->tri_tests += c.tri_tests; o->sphere_tests += c.sphere_tests; o->segments += c.segments; o->bad_samples += c.bad_samples;
}

} // namespace

// ================================================================= C ABI
extern "C" {

const char *orc_last_error(void) { return g_err.c_str(); }

// serde.rs:81-154 (objects/lights) + scene.rs:143-170 (Scene::new)
orc_scene *orc_scene_create(const orc_scene_desc *d) {
    orc_scene *s = new orc_scene();
    for (uint32_t i = 0; i < d->n_models; i++) {
        const orc_model_desc &md = d->models[i];
        Model m;
        m.vertices.resize(md.n_vertices);
        for (uint32_t v = 0; v < md.n_vertices; v++) m.vertices[v] = v3(md.vertices[3 * v], md.vertices[3 * v + 1], md.vertices[3 * v + 2]);
        m.faces.assign(md.faces, md.faces + 3 * (size_t)md.n_faces);
        for (uint32_t f = 0; f < 3 * md.n_faces; f++)
            if (m.faces[f] >= md.n_vertices) { g_err = "face index out of range"; delete s; return nullptr; }
        m.compute_face_normals(); // model.rs:116-128
        if (md.vertex_colours) {
            m.has_vc = true;
            m.vertex_colours.resize(md.n_vertices);
            for (uint32_t v = 0; v < md.n_vertices; v++)
                m.vertex_colours[v] = col(md.vertex_colours[3 * v], md.vertex_colours[3 * v + 1], md.vertex_colours[3 * v + 2]);
        }
        s->models.push_back(std::move(m));
    }
    std::vector<Prim> prims;
    for (uint32_t i = 0; i < d->n_objects; i++) {
        const orc_object_desc &od = d->objects[i];
        Object o;
        o.shape = od.shape;
        o.model = od.model;
        o.smooth = od.smooth_normals != 0;
        // serde.rs:103: Matrix3::rotation(shp.rotation.pitch, shp.rotation.yaw, shp.rotation.roll) -- the first
        // parameter of rotation() is *yaw*, so the YAML pitch drives Ry and the YAML yaw drives Rx.
        o.rot = rotation(od.rot_pitch, od.rot_yaw, od.rot_roll);
        o.material = to_mat(od.material);
        if (od.shape == 0) {
            Prim p;
            p.type = 0;
            p.center = v3(od.center[0], od.center[1], od.center[2]);
            p.radius = od.radius;
            p.index = 0;
            p.v[0] = p.v[1] = p.v[2] = p.n = v3(0, 0, 0);
            prims.push_back(p);
            s->items.push_back(Entity{false, i});
        } else {
            if (od.model < 0 || (uint32_t)od.model >= d->n_models) { g_err = "object model index out of range"; delete s; return nullptr; }
            Model &m = s->models[od.model];
            if (o.smooth) m.compute_vertex_normals(); // serde.rs:115-120
            if (!o.material.fresnel && o.material.m.vertex_albedo && !m.has_vc) { g_err = "Model does not have vertex colours"; delete s; return nullptr; }
            V3 tr = v3(od.translation[0], od.translation[1], od.translation[2]);
            std::vector<Prim> mp = m.resolve_primitives(); // geom.rs:111-117
            for (auto &p : mp) {
                prims.push_back(prim_transform(p, tr, o.rot, od.scale));
                s->items.push_back(Entity{false, i});
            }
        }
        s->objects.push_back(o);
    }
    for (uint32_t i = 0; i < d->n_lights; i++) { // scene.rs:155-164
        const orc_light_desc &ld = d->lights[i];
        Light l{ld.geometry, v3(ld.v[0], ld.v[1], ld.v[2]), ld.radius, col(ld.colour[0], ld.colour[1], ld.colour[2]), ld.intensity};
        s->lights.push_back(l);
        if (ld.geometry == 1) {
            Prim p;
            p.type = 0;
            p.center = l.v;
            p.radius = l.radius;
            p.index = 0;
            p.v[0] = p.v[1] = p.v[2] = p.n = v3(0, 0, 0);
            prims.push_back(p);
            s->items.push_back(Entity{true, i});
        }
    }
    s->sky_type = d->sky.type;
    s->sky_c0 = col(d->sky.c0[0], d->sky.c0[1], d->sky.c0[2]);
    s->sky_c1 = col(d->sky.c1[0], d->sky.c1[1], d->sky.c1[2]);
    s->sky_w = d->sky.width;
    s->sky_h = d->sky.height;
    if (d->sky.type == 2) {
        if (!d->sky.texels || !d->sky.width || !d->sky.height) { g_err = "Hdri sky without texels"; delete s; return nullptr; }
        s->sky_texels.assign(d->sky.texels, d->sky.texels + (size_t)3 * d->sky.width * d->sky.height);
    }
    if (!s->bvh.construct(prims)) { delete s; return nullptr; }
    return s;
}

void orc_scene_destroy(orc_scene *s) { delete s; }
uint32_t orc_scene_num_prims(const orc_scene *s) { return (uint32_t)s->bvh.prims.size(); }
uint32_t orc_scene_num_nodes(const orc_scene *s) { return (uint32_t)s->bvh.next.load(); }
uint32_t orc_scene_bvh_depth(const orc_scene *s) { return s->bvh.depth_of(s->bvh.root); }

int orc_scene_prim_info(const orc_scene *s, uint32_t prim, orc_prim_info *out) {
    if (prim >= s->bvh.prims.size()) return -1;
    const Prim &p = s->bvh.prims[prim];
    const Entity &e = s->items[prim];
    memset(out, 0, sizeof(*out));
    out->type = p.type;
    out->is_light = e.is_light ? 1 : 0;
    out->entity = e.id;
    out->face = (uint32_t)p.index;
    if (p.type == 0) {
        out->v[0] = p.center.x; out->v[1] = p.center.y; out->v[2] = p.center.z; out->v[3] = p.radius;
    } else {
        for (int i = 0; i < 3; i++) { out->v[3 * i] = p.v[i].x; out->v[3 * i + 1] = p.v[i].y; out->v[3 * i + 2] = p.v[i].z; }
        out->normal[0] = p.n.x; out->normal[1] = p.n.y; out->normal[2] = p.n.z;
    }
    AABB bb = prim_aabb(p);
    out->aabb_min[0] = bb.min.x; out->aabb_min[1] = bb.min.y; out->aabb_min[2] = bb.min.z;
    out->aabb_max[0] = bb.max.x; out->aabb_max[1] = bb.max.y; out->aabb_max[2] = bb.max.z;
    return 0;
}

void orc_trace_closest(const orc_scene *s, const double *origins, const double *dirs, size_t n, uint32_t *prim, double *t,
                       double *bary, orc_counters *ctr) {
    unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    unsigned nt = (unsigned)std::min<size_t>(hw, std::max<size_t>(1, n / 4096));
    std::vector<Counters> cs(nt);
    std::vector<std::thread> th;
    for (unsigned w = 0; w < nt; w++) {
        th.emplace_back([&, w] {
            size_t lo = n * w / nt, hi = n * (w + 1) / nt;
            for (size_t i = lo; i < hi; i++) {
                Ray ray = make_ray(v3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), v3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
                Collision c;
                long obj = s->bvh.find_intersection(ray, c, &cs[w]);
                if (obj < 0) {
                    prim[i] = 0xFFFFFFFFu;
                    t[i] = std::numeric_limits<double>::infinity();
                    if (bary) bary[3 * i] = bary[3 * i + 1] = bary[3 * i + 2] = 0.0;
                } else {
                    prim[i] = (uint32_t)obj;
                    t[i] = c.distance;
                    if (bary) { bary[3 * i] = c.has_meta ? c.bx : 0.0; bary[3 * i + 1] = c.has_meta ? c.by : 0.0; bary[3 * i + 2] = c.has_meta ? c.bz : 0.0; }
                }
            }
        });
    }
    for (auto &x : th) x.join();
    Counters tot;
    for (auto &c : cs) tot.add(c);
    tot.extend_rays += n;
    export_counters(tot, ctr);
}

int orc_intersect_prim(const orc_scene *s, uint32_t prim, const double o[3], const double d[3], double *t) {
    if (prim >= s->bvh.prims.size()) return 0;
    Ray ray = make_ray(v3(o[0], o[1], o[2]), v3(d[0], d[1], d[2]));
    Collision c;
    if (!prim_intersect(s->bvh.prims[prim], ray, c, nullptr)) return 0;
    *t = c.distance;
    return 1;
}

void orc_camera_rays(const orc_camera_desc *cd, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h,
                     double *o, double *d, double *weight) {
    Camera cam = to_camera(*cd);
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    size_t i = 0;
    for (uint32_t y = y0; y < y0 + h; y++)
        for (uint32_t x = x0; x < x0 + w; x++, i++) { // worker.rs:170-186 row-major
            Ray ray;
            double wt;
            pixel_sample_ray(cam, key, x, y, sample, ray, wt);
            o[3 * i] = ray.origin.x; o[3 * i + 1] = ray.origin.y; o[3 * i + 2] = ray.origin.z;
            d[3 * i] = ray.direction.x; d[3 * i + 1] = ray.direction.y; d[3 * i + 2] = ray.direction.z;
            weight[i] = wt;
        }
}

void orc_render(const orc_scene *s, const orc_camera_desc *cd, uint64_t seed, uint32_t s0, uint32_t s1, int shard_mode, uint32_t rank,
                uint32_t world, int threads, double *sums, uint32_t *counts, orc_counters *ctr) {
    Camera cam = to_camera(*cd);
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    unsigned nt = threads > 0 ? (unsigned)threads : std::max(1u, std::thread::hardware_concurrency());
    std::atomic<uint32_t> next_row{0};
    std::vector<Counters> cs(nt);
    std::vector<std::thread> th;
    for (unsigned w = 0; w < nt; w++) {
        th.emplace_back([&, w] {
            for (;;) {
                uint32_t y = next_row.fetch_add(1);
                if (y >= cam.height) break;
                for (uint32_t x = 0; x < cam.width; x++) {
                    if (!owns_pixel(shard_mode, rank, world, x, y, cam.width)) continue;
                    size_t ix = (size_t)x + (size_t)y * cam.width; // pixels.rs:49-51
                    for (uint32_t S = s0; S < s1; S++) {
                        Col c = trace_pixel(*s, cam, key, x, y, S, cs[w]);
                        sums[3 * ix] += c.r; sums[3 * ix + 1] += c.g; sums[3 * ix + 2] += c.b; // pixels.rs:19-22
                        counts[ix] += 1;
                    }
                }
            }
        });
    }
    for (auto &x : th) x.join();
    Counters tot;
    for (auto &c : cs) tot.add(c);
    export_counters(tot, ctr);
}

void orc_trace_pixel(const orc_scene *s, const orc_camera_desc *cd, uint64_t seed, uint32_t x, uint32_t y, uint32_t sample, double out[3]) {
    Camera cam = to_camera(*cd);
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    Counters c;
    Col r = trace_pixel(*s, cam, key, x, y, sample, c);
    out[0] = r.r; out[1] = r.g; out[2] = r.b;
}

uint32_t orc_permute(uint32_t i, uint32_t l, uint32_t p) { return cmj_permute(i, l, p); }
double orc_rand_float(uint32_t i, uint32_t p) { return cmj_rand_float(i, p); }
void orc_cmj(uint32_t s, uint32_t m, uint32_t n, uint32_t p, double out[2]) { cmj(s, m, n, p, out[0], out[1]); }
void orc_cmj_disk(uint32_t s, uint32_t m, uint32_t n, uint32_t p, double out[2]) { cmj_disk(s, m, n, p, out[0], out[1]); }
uint64_t orc_morton_code(uint16_t nb, uint16_t x, uint16_t y, uint16_t z) { return morton_code(nb, x, y, z); }
void orc_matmul3(const double a[9], const double b[9], double out[9]) {
    M3 A, B;
    memcpy(A.c, a, sizeof(A.c));
    memcpy(B.c, b, sizeof(B.c));
    M3 C = matmul(A, B);
    memcpy(out, C.c, sizeof(C.c));
}
void orc_rotation(double yaw, double pitch, double roll, double out[9]) {
    M3 r = rotation(yaw, pitch, roll);
    memcpy(out, r.c, sizeof(r.c));
}
void orc_switch_basis(const double v[3], const double normal[3], double out[3]) { // geom.rs:336-349 test body
    V3 i, j, k;
    form_basis(v3(normal[0], normal[1], normal[2]), i, j, k);
    V3 o = switch_basis(v3(v[0], v[1], v[2]), i, j, k);
    out[0] = o.x; out[1] = o.y; out[2] = o.z;
}
void orc_to_bytes(const double *rgb, size_t n, uint8_t *out) { // colour.rs:30-36 + main.rs:114-120
    for (size_t i = 0; i < 3 * n; i++) out[i] = component_to_byte(rgb[i]);
}
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) { philox4x32_10(ctr, key, out); }
void orc_pixel_grid(uint32_t x, uint32_t y, uint32_t w, uint32_t h, uint32_t *out) { // worker.rs:170-186
    for (uint32_t pos = 0; pos < w * h; pos++) {
        out[2 * pos] = x + pos % w;
        out[2 * pos + 1] = y + pos / w;
    }
}
void orc_sample_schedule(uint32_t S, uint32_t out[4]) { sample_schedule(S, out[0], out[1], out[2], out[3]); }
void orc_sky(const orc_scene *s, const double dir[3], double out[3]) {
    Col c = ambient_light(*s, v3(dir[0], dir[1], dir[2]));
    out[0] = c.r; out[1] = c.g; out[2] = c.b;
}
void orc_material_brdf(const orc_material_desc *m, const double vo[3], const double vi[3], const double n[3], double out[3]) {
    Col c = mat_brdf(to_mat(*m), v3(vo[0], vo[1], vo[2]), v3(vi[0], vi[1], vi[2]), v3(n[0], n[1], n[2]));
    out[0] = c.r; out[1] = c.g; out[2] = c.b;
}
int orc_material_sample(const orc_material_desc *m, const double vo[3], const double n[3], const double u[3], double dir[3],
                        double *pdf, double brdf[3]) {
    V3 d;
    Col b;
    bool spec = false;
    if (!mat_sample(to_mat(*m), v3(vo[0], vo[1], vo[2]), v3(n[0], n[1], n[2]), u, d, *pdf, b, spec)) return -1;
    dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
    brdf[0] = b.r; brdf[1] = b.g; brdf[2] = b.b;
    return spec ? 1 : 0;
}
void orc_light_sample(const orc_light_desc *l, const double from[3], double u, double v, double dir[3], double *inv_pdf) {
    Light L{l->geometry, v3(l->v[0], l->v[1], l->v[2]), l->radius, col(l->colour[0], l->colour[1], l->colour[2]), l->intensity};
    V3 d;
    light_sample(L, v3(from[0], from[1], from[2]), u, v, d, *inv_pdf);
    dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}
uint32_t orc_ccrf(uint32_t x) { return (uint32_t)ccrf(x); }

} // extern "C"
