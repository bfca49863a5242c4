// bvh_fast.cpp -- production acceleration structure: binned-SAH binary tree over conservative f32 boxes, leaves of up
// to 4 primitives, emitted as 64-byte two-child nodes over a leaf-ordered primitive stream.
//
// Closest hit does not depend on the shape of the tree (only exact-t ties do), so the production path is free
// to use a better tree than the reference's AAC build (src/bvh.rs:160-343), whose morton scale collapses whenever a scene
// holds a 1e6-radius "plane" sphere (src/bvh.rs:178-182).  One reference behaviour does depend on the boxes and is kept:
// a primitive whose own AABB is flat along an axis can never pass the strict `tmin < tmax` slab test of its leaf
// (src/bvh.rs:16) and is therefore invisible; such primitives are left out of this tree.
#include <cstdio>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <thread>

#include "host_scene.h"

namespace paths {
namespace {

struct FBox {
    float lo[3], hi[3];
    void reset() {
        for (int i = 0; i < 3; i++) {
            lo[i] = std::numeric_limits<float>::infinity();
            hi[i] = -std::numeric_limits<float>::infinity();
        }
    }
    void grow(const FBox &b) {
        for (int i = 0; i < 3; i++) {
            lo[i] = std::min(lo[i], b.lo[i]);
            hi[i] = std::max(hi[i], b.hi[i]);
        }
    }
    float half_area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return dx * dy + dy * dz + dz * dx;
    }
};

inline float round_down(double v) {
    float f = (float)v;
    return ((double)f > v) ? std::nextafterf(f, -std::numeric_limits<float>::infinity()) : f;
}
inline float round_up(double v) {
    float f = (float)v;
    return ((double)f < v) ? std::nextafterf(f, std::numeric_limits<float>::infinity()) : f;
}

struct Ref {
    FBox box;
    float c[3];
    uint32_t prim;
};

struct SNode {
    FBox box;
    int32_t left, right;
    uint32_t first, count; // leaf iff count > 0
};

#ifndef PATHS_SAH_BINS
#define PATHS_SAH_BINS 64 /* with the giants out of the build: 8 / 16 / 64 bins within +-0.7 % of each other, 64 ahead on teapot, bunny and dragon (profiles/r02zz_sah_giants_ab.txt) */
#endif
const int kBins = PATHS_SAH_BINS;
#ifndef PATHS_SAH_SMALL_BINS
#define PATHS_SAH_SMALL_BINS 16 /* bins for nodes of fewer than 512 primitives */
#endif
#ifndef PATHS_SAH_GIANTS
#define PATHS_SAH_GIANTS 1
#endif
uint32_t kMaxLeaf = 4;
float kTraversalCost = 1.0f, kPrimCost = 1.2f;

struct Sah {
    std::vector<Ref> refs;
    std::vector<SNode> nodes; // each task owns a private vector; spliced by the caller

    // builds refs[lo, hi) into `out`, returns local root index
    static int build(std::vector<Ref> &refs, size_t lo, size_t hi, std::vector<SNode> &out, int depth) {
        SNode node;
        node.box.reset();
        FBox cb;
        cb.reset();
        for (size_t i = lo; i < hi; i++) {
            node.box.grow(refs[i].box);
            for (int a = 0; a < 3; a++) {
                cb.lo[a] = std::min(cb.lo[a], refs[i].c[a]);
                cb.hi[a] = std::max(cb.hi[a], refs[i].c[a]);
            }
        }
        size_t count = hi - lo;
        node.left = node.right = -1;
        node.first = (uint32_t)lo;
        node.count = 0;
        int me = (int)out.size();
        out.push_back(node);
        if (count == 1) {
            out[me].count = 1;
            return me;
        }
        // binned SAH over the three axes; small nodes use fewer bins (the per-node cost of 64 bins x 3 axes dominated the build of the
        // 871k-triangle scene: 2.1 s against 0.8 s, for splits that 16 bins find as well once few primitives remain)
        const int nb = count >= 512 ? kBins : std::min(kBins, PATHS_SAH_SMALL_BINS);
        float best_cost = std::numeric_limits<float>::infinity();
        int best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; axis++) {
            float ext = cb.hi[axis] - cb.lo[axis];
            if (!(ext > 0.0f)) continue;
            FBox bb[kBins];
            uint32_t bc[kBins];
            for (int b = 0; b < nb; b++) {
                bb[b].reset();
                bc[b] = 0;
            }
            float k = (float)nb / ext;
            for (size_t i = lo; i < hi; i++) {
                int b = std::min(nb - 1, std::max(0, (int)((refs[i].c[axis] - cb.lo[axis]) * k)));
                bb[b].grow(refs[i].box);
                bc[b]++;
            }
            float right_area[kBins];
            FBox acc;
            acc.reset();
            for (int b = nb - 1; b > 0; b--) {
                acc.grow(bb[b]);
                right_area[b] = acc.half_area();
            }
            acc.reset();
            uint32_t nl = 0;
            for (int b = 0; b < nb - 1; b++) {
                acc.grow(bb[b]);
                nl += bc[b];
                uint32_t nr = (uint32_t)count - nl;
                if (nl == 0 || nr == 0) continue;
                float cost = acc.half_area() * (float)nl + right_area[b + 1] * (float)nr;
                if (cost < best_cost) {
                    best_cost = cost;
                    best_axis = axis;
                    best_bin = b;
                }
            }
        }
        size_t mid;
        if (best_axis >= 0) {
            float leaf_cost = kPrimCost * (float)count;
            float split_cost = kTraversalCost + kPrimCost * best_cost / std::max(node.box.half_area(), 1e-30f);
            if (count <= kMaxLeaf && leaf_cost <= split_cost) {
                out[me].count = (uint32_t)count;
                return me;
            }
            float ext = cb.hi[best_axis] - cb.lo[best_axis];
            float k = (float)nb / ext;
            float clo = cb.lo[best_axis];
            Ref *p = std::partition(refs.data() + lo, refs.data() + hi, [&](const Ref &r) {
                int b = std::min(nb - 1, std::max(0, (int)((r.c[best_axis] - clo) * k)));
                return b <= best_bin;
            });
            mid = (size_t)(p - refs.data());
            if (mid == lo || mid == hi) mid = lo + count / 2;
        } else {
            // all centroids coincide
            if (count <= kMaxLeaf) {
                out[me].count = (uint32_t)count;
                return me;
            }
            mid = lo + count / 2;
        }
        if (depth < 3 && count > 65536) {
            std::vector<SNode> lout, rout;
            int lroot = 0, rroot = 0;
            std::thread tl([&] { lroot = build(refs, lo, mid, lout, depth + 1); });
            rroot = build(refs, mid, hi, rout, depth + 1);
            tl.join();
            int lbase = (int)out.size();
            for (SNode n : lout) {
                if (n.count == 0) {
                    n.left += lbase;
                    n.right += lbase;
                }
                out.push_back(n);
            }
            int rbase = (int)out.size();
            for (SNode n : rout) {
                if (n.count == 0) {
                    n.left += rbase;
                    n.right += rbase;
                }
                out.push_back(n);
            }
            out[me].left = lbase + lroot;
            out[me].right = rbase + rroot;
        } else {
            int l = build(refs, lo, mid, out, depth + 1);
            int r = build(refs, mid, hi, out, depth + 1);
            out[me].left = l;
            out[me].right = r;
        }
        return me;
    }
};

inline int32_t leaf_code(uint32_t first, uint32_t count) { return ~(int32_t)((first << 3) | (count - 1)); }

uint32_t emit2(const std::vector<SNode> &sn, int node, FastTree &out, int32_t &code) {
    const SNode &n = sn[node];
    if (n.count > 0) {
        code = leaf_code(n.first, n.count);
        return 1;
    }
    int32_t me = (int32_t)out.nodes.size();
    out.nodes.push_back(FastNode2());
    code = me;
    int32_t lc, rc;
    uint32_t dl = emit2(sn, n.left, out, lc);
    uint32_t dr = emit2(sn, n.right, out, rc);
    FastNode2 &fn = out.nodes[me];
    const FBox &lb = sn[n.left].box, &rb = sn[n.right].box;
    for (int a = 0; a < 3; a++) {
        fn.c0[2 * a] = lb.lo[a];
        fn.c0[2 * a + 1] = lb.hi[a];
        fn.c1[2 * a] = rb.lo[a];
        fn.c1[2 * a + 1] = rb.hi[a];
    }
    fn.i0 = lc;
    fn.i1 = rc;
    fn.pad[0] = fn.pad[1] = 0;
    return 1 + std::max(dl, dr);
}

// Collapse to four-wide nodes: a node's children are its binary children, with the child of largest surface area replaced by
// its own two children until four slots are used (or only leaves remain).
uint32_t emit4(const std::vector<SNode> &sn, int node, FastTree &out, int32_t &code) {
    const SNode &n = sn[node];
    if (n.count > 0) {
        code = leaf_code(n.first, n.count);
        return 1;
    }
    int kids[4];
    int nk = 0;
    kids[nk++] = n.left;
    kids[nk++] = n.right;
    while (nk < 4) {
        int best = -1;
        float best_area = -1.0f;
        for (int i = 0; i < nk; i++) {
            const SNode &c = sn[kids[i]];
            if (c.count == 0 && c.box.half_area() > best_area) {
                best_area = c.box.half_area();
                best = i;
            }
        }
        if (best < 0) break;
        int open = kids[best];
        kids[best] = sn[open].left;
        kids[nk++] = sn[open].right;
    }
    int32_t me = (int32_t)out.nodes4.size();
    out.nodes4.push_back(FastNode4());
    code = me;
    int32_t codes[4] = {0, 0, 0, 0};
    uint32_t depth = 0;
    for (int i = 0; i < nk; i++) depth = std::max(depth, emit4(sn, kids[i], out, codes[i]));
    FastNode4 &fn = out.nodes4[me];
    std::memset(&fn, 0, sizeof(fn));
    for (int i = 0; i < 4; i++) {
        if (i < nk) {
            const FBox &b = sn[kids[i]].box;
            fn.lox[i] = b.lo[0]; fn.hix[i] = b.hi[0]; fn.loy[i] = b.lo[1]; fn.hiy[i] = b.hi[1]; fn.loz[i] = b.lo[2]; fn.hiz[i] = b.hi[2];
            fn.child[i] = codes[i];
        } else {
            fn.lox[i] = fn.loy[i] = fn.loz[i] = std::numeric_limits<float>::infinity();
            fn.hix[i] = fn.hiy[i] = fn.hiz[i] = -std::numeric_limits<float>::infinity();
            fn.child[i] = 0x7FFFFFFF; // PATHS_CHILD_EMPTY
        }
    }
    return 1 + depth;
}

} // namespace

bool build_fast_tree(const HostScene &s, FastTree &out, std::string &err) {
#ifdef PATHS_DEV_SWITCHES /* development builds only (make EXTRA=-DPATHS_DEV_SWITCHES): a stale variable must not change a release run */
    if (const char *v = getenv("PATHS_SAH_MAXLEAF")) kMaxLeaf = (uint32_t)std::min(8, std::max(1, atoi(v)));
    if (const char *v = getenv("PATHS_SAH_PRIMCOST")) kPrimCost = (float)atof(v);
#endif
    std::vector<Ref> refs;
    refs.reserve(s.prims.size());
    // The slab test evaluates lo * (1/d) - o * (1/d): its error is about one ulp of the ORIGIN's coordinate, wherever the box is.
    // Boxes are therefore grown by a few ulps of the scene's extent so that a ray that hits a triangle is never turned away by
    // that triangle's own box.
    double ext = 1.0;
    if (s.bounds_lo.x <= s.bounds_hi.x)
        ext = std::max({std::fabs(s.bounds_lo.x), std::fabs(s.bounds_hi.x), std::fabs(s.bounds_lo.y), std::fabs(s.bounds_hi.y),
                        std::fabs(s.bounds_lo.z), std::fabs(s.bounds_hi.z), 1.0});
    const float pad = (float)(ext * 4.76837158203125e-7); // 2^-21
    for (size_t i = 0; i < s.prims.size(); i++) {
        const HostPrim &p = s.prims[i];
        if (p.type == 1 && (p.box.lo.x == p.box.hi.x || p.box.lo.y == p.box.hi.y || p.box.lo.z == p.box.hi.z)) continue; // invisible (header)
        Ref r;
        r.box.lo[0] = round_down(p.box.lo.x); r.box.lo[1] = round_down(p.box.lo.y); r.box.lo[2] = round_down(p.box.lo.z);
        r.box.hi[0] = round_up(p.box.hi.x); r.box.hi[1] = round_up(p.box.hi.y); r.box.hi[2] = round_up(p.box.hi.z);
        if (p.type == 1) {
            // the f32 kernel intersects f32-rounded vertices: make the box cover them as well
            const D3 *vs[3] = {&p.a, &p.b, &p.c};
            for (int k = 0; k < 3; k++) {
                float fx = (float)vs[k]->x, fy = (float)vs[k]->y, fz = (float)vs[k]->z;
                r.box.lo[0] = std::min(r.box.lo[0], fx); r.box.hi[0] = std::max(r.box.hi[0], fx);
                r.box.lo[1] = std::min(r.box.lo[1], fy); r.box.hi[1] = std::max(r.box.hi[1], fy);
                r.box.lo[2] = std::min(r.box.lo[2], fz); r.box.hi[2] = std::max(r.box.hi[2], fz);
            }
        }
        if (p.type == 1)
            for (int a = 0; a < 3; a++) {
                r.box.lo[a] -= pad;
                r.box.hi[a] += pad;
            }
        for (int a = 0; a < 3; a++) r.c[a] = 0.5f * r.box.lo[a] + 0.5f * r.box.hi[a];
        r.prim = (uint32_t)i;
        refs.push_back(r);
    }
    out = FastTree();
    if (refs.empty()) {
        // nothing visible: a tree with one empty inner node keeps the kernels branch-free
        FastNode2 fn;
        std::memset(&fn, 0, sizeof(fn));
        for (int a = 0; a < 3; a++) {
            fn.c0[2 * a] = fn.c1[2 * a] = std::numeric_limits<float>::infinity();
            fn.c0[2 * a + 1] = fn.c1[2 * a + 1] = -std::numeric_limits<float>::infinity();
        }
        fn.i0 = fn.i1 = 0;
        out.nodes.push_back(fn);
        out.root = (int32_t)0x80000000; // PATHS_TRAV_DONE: the kernels finish at once
        for (int a = 0; a < 6; a++) out.root_box[a] = (a < 3) ? std::numeric_limits<float>::infinity() : -std::numeric_limits<float>::infinity();
        out.depth = 1;
        FastNode4 f4;
        std::memset(&f4, 0, sizeof(f4));
        for (int i = 0; i < 4; i++) {
            f4.lox[i] = f4.loy[i] = f4.loz[i] = std::numeric_limits<float>::infinity();
            f4.hix[i] = f4.hiy[i] = f4.hiz[i] = -std::numeric_limits<float>::infinity();
        }
        for (int i = 0; i < 4; i++) f4.child[i] = 0x7FFFFFFF;
        out.nodes4.push_back(f4);
        out.root4 = (int32_t)0x80000000;
        out.depth4 = 1;
        (void)err;
        return true;
    }
    // Giant primitives (the scenes' 1e6-radius "plane" spheres: extent above 1024 x the geometric mean extent, the device builder's rule,
    // bvh_device.cu) are kept out of the SAH build and chained above its root.  Every ray is inside their boxes and tests them anyway;
    // left in the build, their centroids stretch the binning domain until all ordinary primitives share one bin, and a giant can end up
    // deep inside the mesh's subtree, where every node on its path inherits its box (bunny.yml, 16 bins: three levels of 3e6-wide nodes
    // inside the mesh; closest-hit time 24.0 -> 20.3 ms per 64 spp once the giants sit at the top, profiles/r02zz_sah_bins_ab.txt).
    size_t n_ord = refs.size();
    if (PATHS_SAH_GIANTS && refs.size() > 1) {
        double log_sum = 0.0;
        std::vector<float> ext_of(refs.size());
        for (size_t i = 0; i < refs.size(); i++) {
            const FBox &b = refs[i].box;
            ext_of[i] = std::max({b.hi[0] - b.lo[0], b.hi[1] - b.lo[1], b.hi[2] - b.lo[2]});
            log_sum += std::log(std::max((double)ext_of[i], 1e-30));
        }
        const double threshold = 1024.0 * std::exp(log_sum / (double)refs.size());
        std::vector<Ref> ord, giants;
        for (size_t i = 0; i < refs.size(); i++) ((double)ext_of[i] > threshold ? giants : ord).push_back(refs[i]);
        if (!giants.empty() && !ord.empty() && giants.size() <= 48) {
            n_ord = ord.size();
            refs = ord;
            refs.insert(refs.end(), giants.begin(), giants.end()); // leaf slots of the giants follow the ordinary primitives'
        }
    }
    std::vector<SNode> sn;
    sn.reserve(refs.size());
    int root = Sah::build(refs, 0, n_ord, sn, 0);
    for (size_t g = n_ord; g < refs.size(); g++) { // chain: new root = {giant leaf, previous root}
        SNode leaf;
        leaf.box = refs[g].box;
        leaf.left = leaf.right = -1;
        leaf.first = (uint32_t)g;
        leaf.count = 1;
        int li = (int)sn.size();
        sn.push_back(leaf);
        SNode up;
        up.box = refs[g].box;
        up.box.grow(sn[root].box);
        up.left = li; // the giant first (equal entry distances go left); searching the ordinary geometry first measured the same
        up.right = root;
        up.first = 0;
        up.count = 0;
        root = (int)sn.size();
        sn.push_back(up);
    }
#ifdef PATHS_DEV_SWITCHES
    if (getenv("PATHS_SAH_REPORT")) { // SAH cost of the binary tree, nodes holding a giant primitive (extent > 1e5) left out
        double inner = 0.0, leaf = 0.0, amax = 0.0;
        size_t n_inner = 0, n_leaf = 0;
        for (const SNode &n : sn) {
            float e = std::max({n.box.hi[0] - n.box.lo[0], n.box.hi[1] - n.box.lo[1], n.box.hi[2] - n.box.lo[2]});
            if (!(e < 1e5f)) continue;
            double a = n.box.half_area();
            amax = std::max(amax, a);
            if (n.count == 0) { inner += a; n_inner++; }
            else { leaf += a * n.count; n_leaf++; }
        }
        if (atoi(getenv("PATHS_SAH_REPORT")) > 1) { // the top of the tree: where the giant primitives sit
            struct Item { int node, depth; };
            std::vector<Item> todo{{root, 0}};
            while (!todo.empty()) {
                Item it = todo.back();
                todo.pop_back();
                const SNode &n = sn[it.node];
                float e = std::max({n.box.hi[0] - n.box.lo[0], n.box.hi[1] - n.box.lo[1], n.box.hi[2] - n.box.lo[2]});
                size_t np = 0;
                std::vector<int> st{it.node};
                while (!st.empty()) { int k = st.back(); st.pop_back(); if (sn[k].count) np += sn[k].count; else { st.push_back(sn[k].left); st.push_back(sn[k].right); } }
                fprintf(stderr, "[top] %*s node %d extent %.4g prims %zu %s\n", 2 * it.depth, "", it.node, e, np, n.count ? "LEAF" : "");
                if (n.count == 0 && it.depth < 5 && (e > 1e5f || it.depth < 2)) { todo.push_back({n.right, it.depth + 1}); todo.push_back({n.left, it.depth + 1}); }
            }
        }
        fprintf(stderr, "[sah] bins %d prims %zu inner %zu leaves %zu area_max %.6g cost_inner %.6g cost_leaf %.6g total(1.0,1.2) %.6g\n", kBins,
                refs.size(), n_inner, n_leaf, amax, inner, leaf, kTraversalCost * inner + kPrimCost * leaf);
    }
#endif
    out.n_giants = (uint32_t)(refs.size() - n_ord);
    out.leaf_prims.resize(refs.size());
    for (size_t i = 0; i < refs.size(); i++) out.leaf_prims[i] = refs[i].prim;
    out.nodes.reserve(refs.size());
    out.depth = emit2(sn, root, out, out.root);
    if (out.depth > 60) { // PATHS_FAST_STACK is 64 entries
        err = "production tree deeper than the traversal stack (" + std::to_string(out.depth) + " levels)";
        return false;
    }
    for (int a = 0; a < 3; a++) {
        out.root_box[a] = sn[root].box.lo[a];
        out.root_box[3 + a] = sn[root].box.hi[a];
    }
    out.nodes4.reserve(refs.size() / 2 + 1);
    out.depth4 = emit4(sn, root, out, out.root4);
    if (out.root4 < 0) { // a single leaf: one node with three empty slots
        FastNode4 f4;
        std::memset(&f4, 0, sizeof(f4));
        for (int i = 0; i < 4; i++) {
            f4.lox[i] = f4.loy[i] = f4.loz[i] = std::numeric_limits<float>::infinity();
            f4.hix[i] = f4.hiy[i] = f4.hiz[i] = -std::numeric_limits<float>::infinity();
        }
        f4.lox[0] = sn[root].box.lo[0]; f4.hix[0] = sn[root].box.hi[0]; f4.loy[0] = sn[root].box.lo[1];
        f4.hiy[0] = sn[root].box.hi[1]; f4.loz[0] = sn[root].box.lo[2]; f4.hiz[0] = sn[root].box.hi[2];
        for (int i = 1; i < 4; i++) f4.child[i] = 0x7FFFFFFF;
        f4.child[0] = out.root4;
        out.nodes4.push_back(f4);
        out.root4 = 0;
        out.depth4 = 2;
    }
    if (out.root < 0) {
        // a single leaf: wrap it in one inner node whose second child is empty
        FastNode2 fn;
        std::memset(&fn, 0, sizeof(fn));
        for (int a = 0; a < 3; a++) {
            fn.c0[2 * a] = sn[root].box.lo[a];
            fn.c0[2 * a + 1] = sn[root].box.hi[a];
            fn.c1[2 * a] = std::numeric_limits<float>::infinity();
            fn.c1[2 * a + 1] = -std::numeric_limits<float>::infinity();
        }
        fn.i0 = out.root;
        fn.i1 = out.root;
        out.nodes.push_back(fn);
        out.root = 0;
        out.depth = 2;
    }
    return true;
}

} // namespace paths
