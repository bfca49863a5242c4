// bvh_ref.cpp -- the reference's acceleration structure, rebuilt deterministically on the host so that the f64
// parity kernels visit nodes in exactly the order BVH::find_intersection does (src/bvh.rs:78-140).
//
// Algorithm: approximate agglomerative clustering as written in src/bvh.rs:160-343 -- morton sort of leaf centres
// (:170-191), recursive bit partition with a median fallback (:208-267), greedy closest-pair merging with the
// reference's "closest" table and its index fix-up pass (:269-325), cost = surface area of the merged box (:345-373).
// The one unspecified step, the tie order of sort_unstable_by_key (:191), is resolved as "input order".
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <limits>
#include <thread>

#include "host_scene.h"

namespace paths {

uint64_t morton_code(uint16_t num_bits, uint16_t x, uint16_t y, uint16_t z) {
    // level l of the code occupies bits [64 - 3*num_bits + 3*l, +3): z, y, x from the low bit (src/bvh.rs:373-384)
    uint64_t code = 0;
    const int base = 64 - 3 * (int)num_bits;
    for (int l = 0; l < (int)num_bits; l++) {
        uint64_t trip = ((uint64_t)((x >> l) & 1) << 2) | ((uint64_t)((y >> l) & 1) << 1) | (uint64_t)((z >> l) & 1);
        code |= trip << (base + 3 * l);
    }
    return code;
}

uint32_t ccrf(uint32_t x) { // src/bvh.rs:149-158, DELTA = 10, EPSILON = 0.01
    double c = std::pow(10.0, 0.5 - 0.01) / 2.0;
    return (uint32_t)std::ceil(c * std::pow((double)x, 0.5 - 0.01));
}

namespace {

const size_t kDelta = 10;

struct ANode {
    Box3 box;
    int32_t left, right; // -1 for leaves
    uint32_t prim;
};

inline uint16_t to_u16(double v) { // Rust `as u16`: saturating, NaN -> 0
    if (!(v > 0.0)) return 0;
    if (v >= 65535.0) return 65535;
    return (uint16_t)v;
}

// The reference forks two threads on each of the first two levels (:223-231).  Merge order inside a subtree never depends
// on its sibling, so the tree is the same however the work is scheduled; here the first three levels fork, nodes are taken
// from a pre-sized arena through an atomic cursor, and emit() renumbers them depth-first so the output is deterministic.
struct Aac {
    std::vector<ANode> arena;
    std::atomic<int> cursor{0};
    int push(const ANode &n) {
        int ix = cursor.fetch_add(1);
        arena[ix] = n;
        return ix;
    }

    inline double area2(int a, int b) const {
        const Box3 &p = arena[a].box, &q = arena[b].box;
        double lx = std::fmin(p.lo.x, q.lo.x), ly = std::fmin(p.lo.y, q.lo.y), lz = std::fmin(p.lo.z, q.lo.z);
        double hx = std::fmax(p.hi.x, q.hi.x), hy = std::fmax(p.hi.y, q.hi.y), hz = std::fmax(p.hi.z, q.hi.z);
        double w = hx - lx, h = hy - ly, d = hz - lz;
        return 2.0 * (w * h + h * d + d * w);
    }
    size_t nearest(const std::vector<int> &cl, size_t ix) const {
        double low = std::numeric_limits<double>::max();
        size_t best = 0;
        const int me = cl[ix];
        for (size_t j = 0; j < cl.size(); j++) {
            if (j == ix) continue;
            double c = area2(me, cl[j]);
            if (c < low) {
                low = c;
                best = j;
            }
        }
        return best;
    }
    // combine_clusters (:269-325)
    void combine(std::vector<int> &cl, size_t n) {
        std::vector<size_t> closest(cl.size());
        for (size_t i = 0; i < cl.size(); i++) closest[i] = nearest(cl, i);
        while (cl.size() > n) {
            double best = std::numeric_limits<double>::max();
            size_t left = 0, right = 0;
            for (size_t i = 0; i < cl.size(); i++) {
                double c = area2(cl[i], cl[closest[i]]);
                if (c < best) {
                    best = c;
                    left = i;
                    right = closest[i];
                }
            }
            if (right < left) std::swap(left, right);
            // the element at the HIGHER position becomes the left child (ClusterNode::new(lc, rc), :295-301)
            int lc = cl[right], rc = cl[left];
            cl.erase(cl.begin() + right);
            cl.erase(cl.begin() + left);
            closest.erase(closest.begin() + right);
            closest.erase(closest.begin() + left);
            ANode nd;
            const Box3 &p = arena[lc].box, &q = arena[rc].box;
            nd.box.lo = d3(std::fmin(p.lo.x, q.lo.x), std::fmin(p.lo.y, q.lo.y), std::fmin(p.lo.z, q.lo.z));
            nd.box.hi = d3(std::fmax(p.hi.x, q.hi.x), std::fmax(p.hi.y, q.hi.y), std::fmax(p.hi.z, q.hi.z));
            nd.left = lc;
            nd.right = rc;
            nd.prim = 0;
            cl.push_back(push(nd));
            closest.push_back(nearest(cl, cl.size() - 1));
            // fix-up pass, including its treatment of the freshly pushed entry (:306-320)
            for (size_t i = 0; i < cl.size(); i++) {
                if (closest[i] == left || closest[i] == right) closest[i] = nearest(cl, i);
                else if (closest[i] >= right) closest[i] -= 2;
                else if (closest[i] >= left) closest[i] -= 1;
            }
        }
    }

    typedef std::pair<int, uint64_t> Item;
    static inline bool bit(uint64_t code, unsigned depth) { return ((code >> (63 - depth)) & 1) != 0; } // :386-388

    // build_tree (:208-238) over items[lo, hi)
    std::vector<int> build(const std::vector<Item> &items, size_t lo, size_t hi, unsigned max_depth, unsigned depth) {
        size_t count = hi - lo;
        std::vector<int> cl;
        if (count < kDelta) {
            for (size_t i = lo; i < hi; i++) cl.push_back(items[i].first);
            combine(cl, ccrf((uint32_t)kDelta));
            return cl;
        }
        size_t split;
        if (depth < max_depth) { // make_partition (:240-267)
            if (bit(items[lo].second, depth)) split = lo;
            else if (!bit(items[hi - 1].second, depth)) split = hi;
            else {
                size_t max0 = 0, min1 = count - 1;
                while (min1 - max0 > 1) {
                    size_t mid = (min1 + max0) / 2;
                    if (bit(items[lo + mid].second, depth)) min1 = mid;
                    else max0 = mid;
                }
                split = lo + min1;
            }
        } else {
            split = lo + count / 2;
        }
        std::vector<int> r;
        if (depth < 3 && count > 4096) {
            std::thread tl([&] { cl = build(items, lo, split, max_depth, depth + 1); });
            r = build(items, split, hi, max_depth, depth + 1);
            tl.join();
        } else {
            cl = build(items, lo, split, max_depth, depth + 1);
            r = build(items, split, hi, max_depth, depth + 1);
        }
        cl.insert(cl.end(), r.begin(), r.end());
        combine(cl, ccrf((uint32_t)count));
        return cl;
    }
};

uint32_t emit(const Aac &a, int node, RefTree &out, int32_t &code, uint32_t depth) {
    // returns subtree depth; code = encoded reference to `node`
    const ANode &n = a.arena[node];
    if (n.left < 0) {
        code = ~(int32_t)n.prim;
        return 1;
    }
    int32_t me = (int32_t)out.nodes.size();
    out.nodes.push_back(RefNode());
    code = me;
    int32_t lcode, rcode;
    uint32_t dl = emit(a, n.left, out, lcode, depth + 1);
    uint32_t dr = emit(a, n.right, out, rcode, depth + 1);
    RefNode &rn = out.nodes[me];
    const Box3 &lb = a.arena[n.left].box, &rb = a.arena[n.right].box;
    rn.lbox[0] = lb.lo.x; rn.lbox[1] = lb.lo.y; rn.lbox[2] = lb.lo.z; rn.lbox[3] = lb.hi.x; rn.lbox[4] = lb.hi.y; rn.lbox[5] = lb.hi.z;
    rn.rbox[0] = rb.lo.x; rn.rbox[1] = rb.lo.y; rn.rbox[2] = rb.lo.z; rn.rbox[3] = rb.hi.x; rn.rbox[4] = rb.hi.y; rn.rbox[5] = rb.hi.z;
    rn.left = lcode;
    rn.right = rcode;
    return 1 + std::max(dl, dr);
}

} // namespace

bool build_ref_tree(const HostScene &s, RefTree &out, std::string &err) {
    const size_t n = s.prims.size();
    if (n == 0) {
        err = "Must have at least one cluster";
        return false;
    }
    Aac a;
    a.arena.resize(2 * n);
    for (size_t i = 0; i < n; i++) { // LeafNode::new (:37-41)
        ANode leaf;
        leaf.box = s.prims[i].box;
        leaf.left = leaf.right = -1;
        leaf.prim = (uint32_t)i;
        a.push(leaf);
    }
    double nb = std::ceil(std::log((double)n) / std::log(4.0)); // f64::log(4.0) (:170)
    if (nb > 16.0) {
        err = "Too many objects to construct BVH"; // :171
        return false;
    }
    unsigned num_bits = (unsigned)to_u16(nb);
    double cap = (double)(1 << num_bits);
    double mx = std::numeric_limits<double>::quiet_NaN();
    std::vector<D3> centre(n);
    for (size_t i = 0; i < n; i++) {
        const Box3 &b = s.prims[i].box;
        centre[i] = (b.lo + b.hi) * 0.5; // src/geom.rs:79-82
        mx = std::fmax(mx, std::fmax(centre[i].x, std::fmax(centre[i].y, centre[i].z)));
    }
    double scale = cap / mx;
    std::vector<Aac::Item> items(n);
    for (size_t i = 0; i < n; i++)
        items[i] = std::make_pair((int)i, morton_code((uint16_t)num_bits, to_u16(centre[i].x * scale), to_u16(centre[i].y * scale),
                                                      to_u16(centre[i].z * scale)));
    std::stable_sort(items.begin(), items.end(), [](const Aac::Item &p, const Aac::Item &q) { return p.second < q.second; });
    std::vector<int> top = a.build(items, 0, n, num_bits, 0);
    a.combine(top, 1);
    int root = top.back();
    out = RefTree();
    out.nodes.reserve(n);
    const Box3 &rb = a.arena[root].box;
    out.root_box[0] = rb.lo.x; out.root_box[1] = rb.lo.y; out.root_box[2] = rb.lo.z;
    out.root_box[3] = rb.hi.x; out.root_box[4] = rb.hi.y; out.root_box[5] = rb.hi.z;
    out.depth = emit(a, root, out, out.root, 0);
    out.n_nodes_total = (uint32_t)a.cursor.load();
    return true;
}

} // namespace paths
