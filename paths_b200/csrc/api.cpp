// api.cpp -- the C ABI of include/paths_b200.h.  No exception crosses this boundary; every failure becomes a
// paths_status plus a thread-local message.  There is no CPU fallback behind any compute entry point.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <exception>
#include <limits>
#include <memory>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include <cuda_runtime_api.h>

#include "loaders.h"
#include "nccl_dyn.h"
#include "scene_device.h"

using namespace paths;

namespace {
thread_local std::string g_err;
paths_status fail(paths_status st, const std::string &msg) {
    g_err = msg;
    return st;
}
double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
} // namespace

struct paths_scene {
    HostScene hs;
    RefTree ref;
    FastTree fast;
    bool has_ref = false, has_fast = false;
    bool device_build = false; // PATHS_BVH_DEVICE: production tree built by bvh_device.cu at staging time
    double ms_ref = 0.0, ms_fast = 0.0;
    DeviceTreeInfo dev_info;   // of the last device build
    uint32_t last_builder = 0;
    struct Staged {
        SceneDevice *sd;
        bool ref, fast;
        bool device_tree;
        DeviceTreeInfo info;
    };
    std::vector<SceneDevice *> retired; // staged copies replaced by a builder switch: renderers may still use them
    std::vector<Staged> staged;
    std::mutex mu;
    ~paths_scene() {
        for (Staged &s : staged) delete s.sd;
        for (SceneDevice *sd : retired) delete sd;
    }
};

struct paths_renderer {
    RendererBase *impl = nullptr;
    paths_scene *scene = nullptr;
    uint32_t w = 0, h = 0;
    uint64_t seed = 0;
    uint32_t rank = 0, world = 1;
    int32_t device = 0;
    paths_render_opts opts; // as resolved at creation (precision AUTO already replaced)
    ~paths_renderer() { delete impl; }
};

struct paths_group {
    std::vector<paths_renderer *> r; // r[i] = rank i on device_ids[i]; r[0] is the combine root
    ~paths_group() {
        for (paths_renderer *x : r) delete x;
    }
};

namespace {

paths_status check_device(int32_t device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(PATHS_ERR_CUDA, std::string("no CUDA device available (") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                                        "); this library has no CPU path");
    if (device < 0 || device >= n) return fail(PATHS_ERR_INVALID_ARG, "device ordinal out of range");
    return PATHS_OK;
}

// stages the scene on `device` with the trees the caller needs; builds a missing tree first
paths_status ensure_staged(paths_scene *s, int device, bool need_ref, bool need_fast, const SceneDevice **out) {
    std::lock_guard<std::mutex> lock(s->mu);
    for (paths_scene::Staged &st : s->staged)
        if (st.sd->device == device && (!need_ref || st.ref) && (!need_fast || st.fast)) {
            *out = st.sd;
            return PATHS_OK;
        }
    std::string err;
    if (need_ref && !s->has_ref) {
        double t0 = now_ms();
        if (!build_ref_tree(s->hs, s->ref, err)) return fail(PATHS_ERR_UNSUPPORTED, err);
        s->ms_ref = now_ms() - t0;
        s->has_ref = true;
    }
    if (need_fast && s->device_build) {
        // the GPU builder works on the prim tables in HBM: stage without a production tree, then build in place
        SceneDevice *sd = upload_scene(device, s->hs, s->has_ref ? &s->ref : nullptr, nullptr, err, true);
        if (!sd) return fail(PATHS_ERR_CUDA, err);
        DeviceTreeInfo info;
        if (build_device_tree(sd, s->hs, info, err)) {
            s->dev_info = info;
            s->last_builder = 1;
            s->staged.push_back(paths_scene::Staged{sd, s->has_ref, true, true, info});
            *out = sd;
            return PATHS_OK;
        }
        delete sd;
        if (err != "small") return fail(PATHS_ERR_CUDA, err);
    }
    if (need_fast && !s->has_fast) {
        double t0 = now_ms();
        if (!build_fast_tree(s->hs, s->fast, err)) return fail(PATHS_ERR_INTERNAL, err);
        s->ms_fast = now_ms() - t0;
        s->has_fast = true;
    }
    // with the device builder selected, a host tree left over from before is only staged when the device build declined ("small")
    const FastTree *ft = (s->has_fast && (!s->device_build || need_fast)) ? &s->fast : nullptr;
    SceneDevice *sd = upload_scene(device, s->hs, s->has_ref ? &s->ref : nullptr, ft, err);
    if (!sd) return fail(PATHS_ERR_CUDA, err);
    if (ft) s->last_builder = 0;
    s->staged.push_back(paths_scene::Staged{sd, s->has_ref, ft != nullptr, false, DeviceTreeInfo()});
    *out = sd;
    return PATHS_OK;
}

// a builder switch: staged copies that carry a production tree are set aside (renderers created earlier keep using them)
void retire_fast_staged(paths_scene *s) {
    std::vector<paths_scene::Staged> keep;
    for (paths_scene::Staged &st : s->staged) {
        if (st.fast) s->retired.push_back(st.sd);
        else keep.push_back(st);
    }
    s->staged.swap(keep);
}

// PATHS_TRACE_* codes name one f32 kernel: (precision F32, kernel code); everything else has kernel 0 (the render's own choice)
int trace_kernel_of(int32_t &precision) {
    if (precision == PATHS_TRACE_CAMERA || precision == PATHS_TRACE_TWO_WIDE || precision == PATHS_TRACE_PLAIN || precision == PATHS_TRACE_FOUR_WIDE) {
        int k = precision;
        precision = PATHS_PRECISION_F32;
        return k;
    }
    return 0;
}

int resolve_precision(const paths_scene *s, int32_t precision) {
    if (precision == PATHS_PRECISION_F32 || precision == PATHS_PRECISION_F64 || precision == PATHS_PRECISION_MIXED) return precision;
    return s->hs.auto_precision;
}

} // namespace

#define PATHS_GUARD_BEGIN try {
#define PATHS_GUARD_END                                                       \
    }                                                                         \
    catch (const std::bad_alloc &) { return fail(PATHS_ERR_INTERNAL, "out of host memory"); } \
    catch (const std::exception &e) { return fail(PATHS_ERR_INTERNAL, e.what()); }            \
    catch (...) { return fail(PATHS_ERR_INTERNAL, "unknown exception"); }

extern "C" {

const char *paths_last_error(void) { return g_err.c_str(); }
const char *paths_version(void) {
#ifdef PATHS_DEBUG_ASSERTS
    return "paths_b200 0.2 (sm_100a) +asserts";
#else
    return "paths_b200 0.2 (sm_100a)";
#endif
}

paths_status paths_device_count(int32_t *count) {
    if (!count) return fail(PATHS_ERR_INVALID_ARG, "null count");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *count = 0;
        return fail(PATHS_ERR_CUDA, cudaGetErrorString(e));
    }
    *count = n;
    return PATHS_OK;
}

paths_status paths_scene_create(const paths_scene_desc *desc, paths_scene **out) {
    PATHS_GUARD_BEGIN
    if (!desc || !out) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    if ((desc->n_models && !desc->models) || (desc->n_objects && !desc->objects) || (desc->n_lights && !desc->lights))
        return fail(PATHS_ERR_INVALID_ARG, "scene description has a count without its array");
    paths_scene *s = new paths_scene();
    std::string err;
    if (!build_host_scene(*desc, s->hs, err)) {
        delete s;
        return fail(PATHS_ERR_INVALID_ARG, err);
    }
    *out = s;
    return PATHS_OK;
    PATHS_GUARD_END
}

paths_status paths_scene_load_yaml(const char *yaml_path, const char *asset_root, uint32_t width, uint32_t height, paths_scene **out,
                                   paths_camera *camera) {
    PATHS_GUARD_BEGIN
    if (!yaml_path || !out || !camera) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    LoadedScene ls;
    std::string err;
    paths_status st = load_scene_yaml(yaml_path, asset_root ? asset_root : "", ls, err);
    if (st != PATHS_OK) return fail(st, err);
    if (width) ls.camera.image_width = width;
    if (height) ls.camera.image_height = height;
    paths_scene_desc desc;
    ls.fill_desc(desc);
    paths_scene *s = new paths_scene();
    if (!build_host_scene(desc, s->hs, err)) {
        delete s;
        return fail(PATHS_ERR_INVALID_ARG, err);
    }
    *camera = ls.camera;
    *out = s;
    return PATHS_OK;
    PATHS_GUARD_END
}

// generate_stress_scene (src/stress.rs:7-66) with a seedable generator: the reference draws from thread_rng, which cannot be seeded.
// Draw order follows the reference: centre x, y, z, radius (random_sphere), then material choice, colour r, g, b and, for Gloss,
// the reflectance (random_material).  SplitMix64; f64 = top 53 bits * 2^-53 (what rand 0.6.5's gen::<f64>() does with its own bits).
paths_status paths_scene_create_stress(uint32_t num_spheres, uint64_t seed, paths_scene **out, paths_camera *camera) {
    PATHS_GUARD_BEGIN
    if (!out || !camera) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    uint64_t state = seed;
    auto next = [&state]() {
        uint64_t z = (state += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    };
    auto unit = [&next]() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); };
    std::vector<paths_object_desc> objs(num_spheres);
    for (paths_object_desc &o : objs) {
        std::memset(&o, 0, sizeof(o));
        o.shape = 0;
        o.smooth_normals = 1;
        o.scale = 1.0;
        o.center[0] = unit() * 100.0 - 50.0;
        o.center[1] = unit() * 100.0 - 50.0;
        o.center[2] = unit() * 100.0;
        o.radius = unit() * 5.0;
        int choice = (int)(unit() * 3.0); // gen_range(0, 3)
        double colour[3] = {unit(), unit(), unit()};
        paths_basic_material &m = o.material.m;
        for (int k = 0; k < 3; k++) m.albedo[k] = colour[k];
        if (choice == 0) {
            m.type = PATHS_MAT_GLOSS;
            m.reflectance = 1.0 + unit() * 2.0;
            m.metalness = 0.0;
        } else if (choice == 1) m.type = PATHS_MAT_LAMBERTIAN;
        else {
            m.type = PATHS_MAT_MIRROR;
            for (int k = 0; k < 3; k++) m.albedo[k] = 0.0;
        }
    }
    paths_scene_desc desc;
    std::memset(&desc, 0, sizeof(desc));
    desc.n_objects = num_spheres;
    desc.objects = objs.data();
    desc.sky.type = 0; // Flat 0.8 grey, no lights
    for (int k = 0; k < 3; k++) desc.sky.c0[k] = 0.8;
    paths_camera cam;
    std::memset(&cam, 0, sizeof(cam));
    cam.image_width = 720;
    cam.image_height = 480;
    cam.location[0] = 0.0; cam.location[1] = -5.0; cam.location[2] = -13.0;
    cam.pitch = 0.0; cam.yaw = 0.0; cam.roll = -0.3;
    cam.sensor_width = 0.036; cam.sensor_height = 0.024; cam.focal_length = 0.05; cam.focus_distance = 10.0; cam.aperture = 8.0;
    paths_status st = paths_scene_create(&desc, out);
    if (st == PATHS_OK) *camera = cam;
    return st;
    PATHS_GUARD_END
}

void paths_scene_destroy(paths_scene *s) { delete s; }

paths_status paths_scene_build_bvh(paths_scene *s, int32_t which) {
    PATHS_GUARD_BEGIN
    if (!s) return fail(PATHS_ERR_INVALID_ARG, "null scene");
    std::lock_guard<std::mutex> lock(s->mu);
    std::string err;
    if (which == PATHS_BVH_REFERENCE) {
        if (!s->has_ref) {
            double t0 = now_ms();
            if (!build_ref_tree(s->hs, s->ref, err)) return fail(PATHS_ERR_UNSUPPORTED, err);
            s->ms_ref = now_ms() - t0;
            s->has_ref = true;
        }
    } else if (which == PATHS_BVH_FAST) {
        if (!s->has_fast) {
            double t0 = now_ms();
            if (!build_fast_tree(s->hs, s->fast, err)) return fail(PATHS_ERR_INTERNAL, err);
            s->ms_fast = now_ms() - t0;
            s->has_fast = true;
        }
        if (s->device_build) retire_fast_staged(s);
        s->device_build = false;
    } else if (which == PATHS_BVH_DEVICE) {
        if (!s->device_build) retire_fast_staged(s);
        s->device_build = true;
    } else return fail(PATHS_ERR_INVALID_ARG, "unknown tree kind");
    return PATHS_OK;
    PATHS_GUARD_END
}

} // extern "C" (reopened below)

namespace {
struct TreeCheck {
    const HostScene &hs;
    const std::vector<float> &leaves;
    std::vector<uint32_t> seen;
    uint32_t bad = 0;
    size_t visited = 0;
    explicit TreeCheck(const HostScene &h, const std::vector<float> &l) : hs(h), leaves(l), seen(h.prims.size(), 0) {}
    static bool inside(const float *ilo, const float *ihi, const float *olo, const float *ohi) {
        for (int a = 0; a < 3; a++)
            if (!(ilo[a] >= olo[a] && ihi[a] <= ohi[a])) return false;
        return true;
    }
    // every primitive of leaf `code` lies inside (lo, hi)
    void leaf(int32_t code, const float *lo, const float *hi) {
        uint32_t c = (uint32_t)~code, first = c >> 3, cnt = (c & 7u) + 1u;
        for (uint32_t k = first; k < first + cnt; k++) {
            if (12 * (size_t)k + 12 > leaves.size()) {
                bad++;
                continue;
            }
            uint32_t tag;
            std::memcpy(&tag, &leaves[12 * (size_t)k + 11], 4);
            uint32_t prim = tag & 0x7FFFFFFFu;
            if (prim >= hs.prims.size()) {
                bad++;
                continue;
            }
            seen[prim]++;
            const HostPrim &p = hs.prims[prim];
            float plo[3], phi[3];
            if (p.type == 1) {
                const D3 *v[3] = {&p.a, &p.b, &p.c};
                for (int a = 0; a < 3; a++) {
                    plo[a] = std::numeric_limits<float>::infinity();
                    phi[a] = -plo[a];
                }
                for (int j = 0; j < 3; j++) {
                    float f[3] = {(float)v[j]->x, (float)v[j]->y, (float)v[j]->z};
                    for (int a = 0; a < 3; a++) {
                        plo[a] = std::min(plo[a], f[a]);
                        phi[a] = std::max(phi[a], f[a]);
                    }
                }
            } else { // the box of a sphere was rounded outward from f64: compare in f64
                double c3[3] = {p.a.x, p.a.y, p.a.z};
                bool ok = true;
                for (int a = 0; a < 3; a++) ok = ok && (double)lo[a] <= c3[a] - p.b.x && (double)hi[a] >= c3[a] + p.b.x;
                if (!ok) bad++;
                continue;
            }
            if (!inside(plo, phi, lo, hi)) bad++;
        }
    }
    void finish() {
        for (size_t i = 0; i < hs.prims.size(); i++) {
            const HostPrim &p = hs.prims[i];
            bool hidden = p.type == 1 && (p.box.lo.x == p.box.hi.x || p.box.lo.y == p.box.hi.y || p.box.lo.z == p.box.hi.z);
            if (seen[i] != (hidden ? 0u : 1u)) bad++;
        }
    }
};

// two-wide: 16 floats per node {c0 lo.x hi.x lo.y hi.y | c1 ... | c0 lo.z hi.z c1 lo.z hi.z | codes}
uint32_t check_tree2(const HostScene &hs, const std::vector<float> &nodes, const std::vector<float> &leaves, int32_t root) {
    TreeCheck tc(hs, leaves);
    if (root == (int32_t)0x80000000) {
        tc.finish();
        return tc.bad;
    }
    struct Item {
        int32_t node;
        float lo[3], hi[3];
        bool has_box;
    };
    std::vector<Item> stack;
    Item r;
    r.node = root;
    r.has_box = false;
    stack.push_back(r);
    while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        if (it.node < 0 || 16 * (size_t)it.node + 16 > nodes.size() || ++tc.visited > nodes.size() / 16 + 1) {
            tc.bad++;
            continue;
        }
        const float *n = &nodes[16 * (size_t)it.node];
        int32_t codes[2];
        std::memcpy(codes, n + 12, 8);
        for (int c = 0; c < 2; c++) {
            Item ch;
            ch.lo[0] = n[4 * c + 0]; ch.hi[0] = n[4 * c + 1]; ch.lo[1] = n[4 * c + 2]; ch.hi[1] = n[4 * c + 3];
            ch.lo[2] = n[8 + 2 * c]; ch.hi[2] = n[9 + 2 * c];
            if (!(ch.lo[0] <= ch.hi[0])) continue; // empty slot (single-leaf wrapper)
            if (it.has_box && !TreeCheck::inside(ch.lo, ch.hi, it.lo, it.hi)) tc.bad++;
            if (c == 1 && codes[1] == codes[0] && codes[0] < 0) continue; // wrapper repeats its only leaf
            if (codes[c] < 0) tc.leaf(codes[c], ch.lo, ch.hi);
            else {
                ch.node = codes[c];
                ch.has_box = true;
                stack.push_back(ch);
            }
        }
    }
    tc.finish();
    return tc.bad;
}

// four-wide: 32 floats per node {lox[4] hix[4] loy[4] hiy[4] loz[4] hiz[4] codes[4] pad[4]}
uint32_t check_tree4(const HostScene &hs, const std::vector<float> &nodes, const std::vector<float> &leaves, int32_t root) {
    TreeCheck tc(hs, leaves);
    if (root == (int32_t)0x80000000) {
        tc.finish();
        return tc.bad;
    }
    struct Item {
        int32_t node;
        float lo[3], hi[3];
        bool has_box;
    };
    std::vector<Item> stack;
    Item r;
    r.node = root;
    r.has_box = false;
    stack.push_back(r);
    while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        if (it.node < 0 || 32 * (size_t)it.node + 32 > nodes.size() || ++tc.visited > nodes.size() / 32 + 1) {
            tc.bad++;
            continue;
        }
        const float *n = &nodes[32 * (size_t)it.node];
        int32_t codes[4];
        std::memcpy(codes, n + 24, 16);
        for (int c = 0; c < 4; c++) {
            if (codes[c] == 0x7FFFFFFF) continue;
            Item ch;
            ch.lo[0] = n[c]; ch.hi[0] = n[4 + c]; ch.lo[1] = n[8 + c]; ch.hi[1] = n[12 + c]; ch.lo[2] = n[16 + c]; ch.hi[2] = n[20 + c];
            if (it.has_box && !TreeCheck::inside(ch.lo, ch.hi, it.lo, it.hi)) tc.bad++;
            if (codes[c] < 0) tc.leaf(codes[c], ch.lo, ch.hi);
            else {
                ch.node = codes[c];
                ch.has_box = true;
                stack.push_back(ch);
            }
        }
    }
    tc.finish();
    return tc.bad;
}
} // namespace

paths_status paths_scene_check_bvh(paths_scene *s, int32_t device, uint32_t *violations) {
    PATHS_GUARD_BEGIN
    if (!s || !violations) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    paths_status st = check_device(device);
    if (st != PATHS_OK) return st;
    const SceneDevice *sd = nullptr;
    st = ensure_staged(s, device, false, true, &sd);
    if (st != PATHS_OK) return st;
    DeviceTreeInfo info;
    {
        std::lock_guard<std::mutex> lock(s->mu);
        for (paths_scene::Staged &e : s->staged)
            if (e.sd == sd) {
                if (e.device_tree) info = e.info;
                else {
                    info.nodes2 = (uint32_t)s->fast.nodes.size();
                    info.nodes4 = (uint32_t)s->fast.nodes4.size();
                    info.leaves = (uint32_t)s->fast.leaf_prims.size();
                }
            }
    }
    std::vector<float> n2, n4, lv;
    std::string err;
    if (!download_fast_tree(sd, info, n2, n4, lv, err)) return fail(PATHS_ERR_CUDA, err);
    *violations = check_tree2(s->hs, n2, lv, sd->view.fast_root) + check_tree4(s->hs, n4, lv, sd->view.fast4_root);
    return PATHS_OK;
    PATHS_GUARD_END
}

extern "C" {

paths_status paths_scene_get_info(const paths_scene *s, paths_scene_info *info) {
    if (!s || !info) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    std::memset(info, 0, sizeof(*info));
    info->n_prims = (uint32_t)s->hs.prims.size();
    info->n_triangles = s->hs.n_tris;
    info->n_spheres = s->hs.n_spheres;
    info->n_objects = (uint32_t)s->hs.objects.size();
    info->n_lights = (uint32_t)s->hs.lights.size();
    if (s->has_ref) {
        info->ref_nodes = s->ref.n_nodes_total;
        info->ref_depth = s->ref.depth;
    }
    if (s->has_fast) {
        info->fast_nodes = (uint32_t)s->fast.nodes.size();
        info->fast_depth = s->fast.depth;
    }
    info->hidden_prims = s->hs.hidden;
    info->auto_precision = s->hs.auto_precision;
    info->bounds_min[0] = s->hs.bounds_lo.x; info->bounds_min[1] = s->hs.bounds_lo.y; info->bounds_min[2] = s->hs.bounds_lo.z;
    info->bounds_max[0] = s->hs.bounds_hi.x; info->bounds_max[1] = s->hs.bounds_hi.y; info->bounds_max[2] = s->hs.bounds_hi.z;
    info->build_ms_ref = s->ms_ref;
    info->build_ms_fast = s->ms_fast;
    info->build_ms_device = s->dev_info.ms;
    info->bvh_builder = s->last_builder;
    info->device_nodes2 = s->dev_info.nodes2;
    info->device_nodes4 = s->dev_info.nodes4;
    info->device_depth2 = s->dev_info.depth2;
    info->device_depth4 = s->dev_info.depth4;
    info->giant_prims = (s->last_builder == 0 && s->has_fast) ? s->fast.n_giants : s->dev_info.giants;
    return PATHS_OK;
}

paths_status paths_scene_prim_entity(const paths_scene *s, uint32_t prim, int32_t *is_light, uint32_t *entity, uint32_t *face) {
    if (!s) return fail(PATHS_ERR_INVALID_ARG, "null scene");
    if (prim >= s->hs.prims.size()) return fail(PATHS_ERR_INVALID_ARG, "prim id out of range");
    const HostPrim &p = s->hs.prims[prim];
    if (is_light) *is_light = p.is_light;
    if (entity) *entity = p.entity;
    if (face) *face = p.face;
    return PATHS_OK;
}

paths_status paths_scene_prim_geometry(const paths_scene *s, uint32_t prim, double out[12], int32_t *is_triangle) {
    if (!s || !out) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    if (prim >= s->hs.prims.size()) return fail(PATHS_ERR_INVALID_ARG, "prim id out of range");
    const HostPrim &p = s->hs.prims[prim];
    for (int i = 0; i < 12; i++) out[i] = 0.0;
    if (p.type == 1) {
        const D3 *v[4] = {&p.a, &p.b, &p.c, &p.n};
        for (int k = 0; k < 4; k++) {
            out[3 * k] = v[k]->x;
            out[3 * k + 1] = v[k]->y;
            out[3 * k + 2] = v[k]->z;
        }
    } else {
        out[0] = p.a.x; out[1] = p.a.y; out[2] = p.a.z; out[3] = p.b.x;
    }
    if (is_triangle) *is_triangle = p.type == 1;
    return PATHS_OK;
}

paths_status paths_trace_closest(paths_scene *s, int32_t device, const double *origins, const double *dirs, size_t n, uint32_t *prim_id, double *t,
                                 double *bary, int32_t precision) {
    PATHS_GUARD_BEGIN
    if (!s || (n && (!origins || !dirs || !prim_id || !t))) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    paths_status st = check_device(device);
    if (st != PATHS_OK) return st;
    const int kernel = trace_kernel_of(precision);
    int prec = resolve_precision(s, precision);
    const SceneDevice *sd = nullptr;
    st = ensure_staged(s, device, prec == PATHS_PRECISION_F64, prec != PATHS_PRECISION_F64, &sd);
    if (st != PATHS_OK) return st;
    std::string err;
    st = prec == PATHS_PRECISION_F64     ? trace_closest_f64(sd, origins, dirs, n, prim_id, t, bary, err)
         : prec == PATHS_PRECISION_MIXED ? trace_closest_mixed(sd, origins, dirs, n, prim_id, t, bary, err)
                                         : trace_closest_f32(sd, origins, dirs, n, prim_id, t, bary, kernel, err);
    return st == PATHS_OK ? st : fail(st, err);
    PATHS_GUARD_END
}

// No counterpart in the reference: the f32 per-ray seam with its scratch in a static
// device array.  Nothing on this path allocates: a scene that is not staged on the device yet is refused instead of staged.
paths_status paths_trace_closest_static(paths_scene *s, int32_t device, const double *origins, const double *dirs, size_t n, uint32_t *prim_id,
                                        double *t, double *bary) {
    PATHS_GUARD_BEGIN
    if (!s || (n && (!origins || !dirs || !prim_id || !t))) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    paths_status st = check_device(device);
    if (st != PATHS_OK) return st;
    const SceneDevice *sd = nullptr;
    {
        std::lock_guard<std::mutex> lock(s->mu);
        for (paths_scene::Staged &e : s->staged)
            if (!sd && e.sd->device == device && e.fast) sd = e.sd; // the copy ensure_staged would hand paths_trace_closest
    }
    if (!sd)
        return fail(PATHS_ERR_UNSUPPORTED, "paths_trace_closest_static: the scene is not staged on this device yet (staging allocates; call "
                                           "paths_trace_closest, paths_scene_check_bvh or paths_renderer_create on it once first)");
    std::string err;
    st = trace_closest_static_f32(sd, origins, dirs, n, prim_id, t, bary, err);
    return st == PATHS_OK ? st : fail(st, err);
    PATHS_GUARD_END
}

paths_status paths_trace_occluded(paths_scene *s, int32_t device, const double *origins, const double *dirs, const uint32_t *light, size_t n,
                                  uint8_t *occluded, int32_t precision) {
    PATHS_GUARD_BEGIN
    if (!s || (n && (!origins || !dirs || !occluded))) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    if (light)
        for (size_t i = 0; i < n; i++)
            if (light[i] >= s->hs.lights.size()) return fail(PATHS_ERR_INVALID_ARG, "light index out of range");
    paths_status st = check_device(device);
    if (st != PATHS_OK) return st;
    const int kernel = trace_kernel_of(precision);
    if (kernel == PATHS_TRACE_CAMERA) return fail(PATHS_ERR_INVALID_ARG, "PATHS_TRACE_CAMERA names a closest-hit routine; visibility rays never run it");
    int prec = resolve_precision(s, precision);
    const SceneDevice *sd = nullptr;
    st = ensure_staged(s, device, prec == PATHS_PRECISION_F64, prec != PATHS_PRECISION_F64, &sd);
    if (st != PATHS_OK) return st;
    std::string err;
    st = prec == PATHS_PRECISION_F64     ? trace_occluded_f64(sd, origins, dirs, light, n, occluded, err)
         : prec == PATHS_PRECISION_MIXED ? trace_occluded_mixed(sd, origins, dirs, light, n, occluded, err)
                                         : trace_occluded_f32(sd, origins, dirs, light, n, occluded, kernel, err);
    return st == PATHS_OK ? st : fail(st, err);
    PATHS_GUARD_END
}

void paths_render_opts_default(paths_render_opts *o) {
    if (!o) return;
    std::memset(o, 0, sizeof(*o));
    o->precision = PATHS_PRECISION_AUTO;
    o->shard_mode = PATHS_SHARD_SAMPLES;
    o->world = 1;
    o->seed = 1;
    o->max_bounces = 10; // trace.rs:14
}

paths_status paths_renderer_create(paths_scene *s, const paths_camera *cam, const paths_render_opts *opts_in, paths_renderer **out) {
    PATHS_GUARD_BEGIN
    if (!s || !cam || !out) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    paths_render_opts opts;
    if (opts_in) opts = *opts_in;
    else paths_render_opts_default(&opts);
    if (cam->image_width == 0 || cam->image_height == 0 || (uint64_t)cam->image_width * cam->image_height > 0x40000000ull)
        return fail(PATHS_ERR_INVALID_ARG, "bad image size");
    if (opts.max_bounces > 250) return fail(PATHS_ERR_INVALID_ARG, "max_bounces > 250");
    if (opts.shard_mode != PATHS_SHARD_SAMPLES && opts.shard_mode != PATHS_SHARD_TILES) return fail(PATHS_ERR_INVALID_ARG, "unknown shard mode");
    paths_status st = check_device(opts.device);
    if (st != PATHS_OK) return st;
    int prec = resolve_precision(s, opts.precision);
    opts.precision = prec;
    const SceneDevice *sd = nullptr;
    st = ensure_staged(s, opts.device, prec == PATHS_PRECISION_F64, prec != PATHS_PRECISION_F64, &sd);
    if (st != PATHS_OK) return st;
    std::string err;
    RendererBase *impl = prec == PATHS_PRECISION_F32 ? make_renderer_f32(sd, *cam, opts, err) : make_renderer_f64(sd, *cam, opts, err);
    if (!impl) return fail(PATHS_ERR_CUDA, err);
    paths_renderer *r = new paths_renderer();
    r->impl = impl;
    r->scene = s;
    r->w = cam->image_width;
    r->h = cam->image_height;
    r->seed = opts.seed;
    r->rank = opts.rank;
    r->world = std::max<uint32_t>(opts.world, 1);
    r->device = opts.device;
    r->opts = opts;
    *out = r;
    return PATHS_OK;
    PATHS_GUARD_END
}

// ------------------------------------------------------------------ frame combine over NCCL
const char *paths_nccl_info(void) {
    static thread_local std::string info;
    std::string err;
    const NcclApi *api = nccl_api(err);
    if (!api) info = "NCCL unavailable: " + err;
    else {
        int v = 0;
        api->GetVersion(&v);
        info = "NCCL " + std::to_string(v / 10000) + "." + std::to_string(v / 100 % 100) + "." + std::to_string(v % 100) + " from " +
               (api->origin.empty() ? std::string("the process image") : api->origin);
    }
    return info.c_str();
}
paths_status paths_comm_unique_id(uint8_t id[PATHS_COMM_ID_BYTES]) {
    PATHS_GUARD_BEGIN
    static_assert(PATHS_COMM_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "ncclUniqueId size");
    if (!id) return fail(PATHS_ERR_INVALID_ARG, "null id");
    std::string err;
    const NcclApi *api = nccl_api(err);
    if (!api) return fail(PATHS_ERR_NCCL, err);
    ncclUniqueId uid;
    ncclResult_t rr = api->GetUniqueId(&uid);
    if (rr != ncclSuccess) return fail(PATHS_ERR_NCCL, std::string("ncclGetUniqueId: ") + api->GetErrorString(rr));
    std::memcpy(id, uid.internal, PATHS_COMM_ID_BYTES);
    return PATHS_OK;
    PATHS_GUARD_END
}

paths_status paths_renderer_attach_comm(paths_renderer *r, const uint8_t id[PATHS_COMM_ID_BYTES]) {
    PATHS_GUARD_BEGIN
    if (!r || !id) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    if (r->impl->has_comm()) return fail(PATHS_ERR_INVALID_ARG, "a communicator is already attached to this renderer");
    if (r->world < 2) return PATHS_OK; // an unsharded renderer has nothing to combine
    std::string err;
    const NcclApi *api = nccl_api(err);
    if (!api) return fail(PATHS_ERR_NCCL, err);
    if (cudaSetDevice(r->device) != cudaSuccess) return fail(PATHS_ERR_CUDA, "cudaSetDevice failed");
    ncclUniqueId uid;
    std::memcpy(uid.internal, id, PATHS_COMM_ID_BYTES);
    ncclComm_t comm = nullptr;
    ncclResult_t rr = api->CommInitRank(&comm, (int)r->world, uid, (int)r->rank);
    if (rr != ncclSuccess) return fail(PATHS_ERR_NCCL, std::string("ncclCommInitRank: ") + api->GetErrorString(rr));
    paths_status st = r->impl->attach_comm(comm, true);
    return st == PATHS_OK ? st : fail(st, r->impl->err);
    PATHS_GUARD_END
}

paths_status paths_group_create(paths_scene *s, const paths_camera *cam, const paths_render_opts *opts_in, int32_t n_devices,
                                const int32_t *device_ids, paths_group **out) {
    PATHS_GUARD_BEGIN
    if (!s || !cam || !out) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    if (n_devices < 1 || n_devices > 64) return fail(PATHS_ERR_INVALID_ARG, "n_devices out of range");
    paths_render_opts opts;
    if (opts_in) opts = *opts_in;
    else paths_render_opts_default(&opts);
    std::vector<int> devs(n_devices);
    for (int i = 0; i < n_devices; i++) {
        devs[i] = device_ids ? device_ids[i] : i;
        paths_status st = check_device(devs[i]);
        if (st != PATHS_OK) return st;
        for (int j = 0; j < i; j++)
            if (devs[j] == devs[i]) return fail(PATHS_ERR_INVALID_ARG, "a device is listed twice");
    }
    std::unique_ptr<paths_group> g(new paths_group());
    for (int i = 0; i < n_devices; i++) {
        paths_render_opts o = opts;
        o.device = devs[i];
        o.rank = (uint32_t)i;
        o.world = (uint32_t)n_devices;
        paths_renderer *r = nullptr;
        paths_status st = paths_renderer_create(s, cam, &o, &r);
        if (st != PATHS_OK) return st;
        g->r.push_back(r);
    }
    if (n_devices > 1) {
        std::string err;
        const NcclApi *api = nccl_api(err);
        if (!api) return fail(PATHS_ERR_NCCL, err);
        std::vector<ncclComm_t> comms(n_devices, nullptr);
        ncclResult_t rr = api->CommInitAll(comms.data(), n_devices, devs.data());
        if (rr != ncclSuccess) return fail(PATHS_ERR_NCCL, std::string("ncclCommInitAll: ") + api->GetErrorString(rr));
        for (int i = 0; i < n_devices; i++) g->r[i]->impl->attach_comm(comms[i], true);
    }
    *out = g.release();
    return PATHS_OK;
    PATHS_GUARD_END
}

paths_status paths_group_render(paths_group *g, uint32_t spp) {
    PATHS_GUARD_BEGIN
    if (!g) return fail(PATHS_ERR_INVALID_ARG, "null group");
    // one host thread drives every device: all launches are asynchronous, and the combine of all ranks is issued inside one NCCL
    // group, the documented pattern for several communicators in one thread
    for (paths_renderer *r : g->r) {
        paths_status st = r->impl->render(spp, false);
        if (st != PATHS_OK) return fail(st, r->impl->err);
    }
    if (g->r.size() > 1) {
        std::string err;
        const NcclApi *api = nccl_api(err);
        if (!api) return fail(PATHS_ERR_NCCL, err);
        api->GroupStart();
        paths_status bad = PATHS_OK;
        std::string why;
        for (paths_renderer *r : g->r) {
            paths_status st = r->impl->combine_post(0);
            if (st != PATHS_OK && bad == PATHS_OK) {
                bad = st;
                why = r->impl->err;
            }
        }
        ncclResult_t rr = api->GroupEnd(); // the grouped sends / receives / reduces are enqueued on the ranks' streams here
        if (bad != PATHS_OK) return fail(bad, why);
        if (rr != ncclSuccess) return fail(PATHS_ERR_NCCL, std::string("ncclGroupEnd: ") + api->GetErrorString(rr));
        for (paths_renderer *r : g->r) {
            paths_status st = r->impl->combine_finish(0);
            if (st != PATHS_OK) return fail(st, r->impl->err);
        }
    }
    for (paths_renderer *r : g->r) {
        paths_status st = r->impl->sync();
        if (st != PATHS_OK) return fail(st, r->impl->err);
    }
    return PATHS_OK;
    PATHS_GUARD_END
}

#define PATHS_GROUP_ROOT(expr)                                            \
    PATHS_GUARD_BEGIN                                                     \
    if (!g) return fail(PATHS_ERR_INVALID_ARG, "null group");             \
    paths_renderer *r = g->r[0];                                          \
    paths_status st = (expr);                                             \
    return st == PATHS_OK ? st : fail(st, r->impl->err);                  \
    PATHS_GUARD_END

paths_status paths_group_image_f32(paths_group *g, float *rgb) {
    if (!rgb) return fail(PATHS_ERR_INVALID_ARG, "null buffer");
    PATHS_GROUP_ROOT(r->impl->image_f32(rgb))
}
paths_status paths_group_image_rgb24(paths_group *g, uint8_t *rgb) {
    if (!rgb) return fail(PATHS_ERR_INVALID_ARG, "null buffer");
    PATHS_GROUP_ROOT(r->impl->image_rgb24(rgb))
}
paths_status paths_group_get_sums(paths_group *g, double *sums4) {
    if (!sums4) return fail(PATHS_ERR_INVALID_ARG, "null buffer");
    PATHS_GROUP_ROOT(r->impl->get_sums(sums4))
}
paths_status paths_group_get_stats(paths_group *g, paths_stats *out) {
    PATHS_GUARD_BEGIN
    if (!g || !out) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    paths_stats tot;
    std::memset(&tot, 0, sizeof(tot));
    for (size_t i = 0; i < g->r.size(); i++) {
        paths_stats st;
        paths_status rc = g->r[i]->impl->get_stats(&st);
        if (rc != PATHS_OK) return fail(rc, g->r[i]->impl->err);
        if (i == 0) tot = st;
        else {
            tot.samples += st.samples; tot.extend_rays += st.extend_rays; tot.shadow_rays += st.shadow_rays; tot.bad_samples += st.bad_samples;
            tot.kernel_launches += st.kernel_launches; tot.passes += st.passes; tot.stack_overflows += st.stack_overflows; tot.assert_failures += st.assert_failures;
            tot.n_extend_launches += st.n_extend_launches; tot.n_shadow_launches += st.n_shadow_launches;
            tot.ms_raygen += st.ms_raygen; tot.ms_extend += st.ms_extend; tot.ms_shade += st.ms_shade; tot.ms_shadow += st.ms_shadow;
            tot.ms_accumulate += st.ms_accumulate; tot.ms_total += st.ms_total;
            tot.ms_device = std::max(tot.ms_device, st.ms_device);
        }
    }
    *out = tot;
    return PATHS_OK;
    PATHS_GUARD_END
}
paths_status paths_group_renderer(paths_group *g, int32_t i, paths_renderer **out) {
    if (!g || !out) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    if (i < 0 || (size_t)i >= g->r.size()) return fail(PATHS_ERR_INVALID_ARG, "rank out of range");
    *out = g->r[i];
    return PATHS_OK;
}
paths_status paths_group_set_camera(paths_group *g, const double location[3], const double rot[9]) {
    PATHS_GUARD_BEGIN
    if (!g || !location || !rot) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    for (paths_renderer *r : g->r) {
        paths_status st = r->impl->set_camera(location, rot);
        if (st != PATHS_OK) return fail(st, r->impl->err);
    }
    return PATHS_OK;
    PATHS_GUARD_END
}
paths_status paths_group_reset(paths_group *g) {
    PATHS_GUARD_BEGIN
    if (!g) return fail(PATHS_ERR_INVALID_ARG, "null group");
    for (paths_renderer *r : g->r) {
        paths_status st = r->impl->reset();
        if (st != PATHS_OK) return fail(st, r->impl->err);
    }
    return PATHS_OK;
    PATHS_GUARD_END
}
void paths_group_destroy(paths_group *g) { delete g; }

#define PATHS_FORWARD(expr)                                         \
    PATHS_GUARD_BEGIN                                               \
    if (!r) return fail(PATHS_ERR_INVALID_ARG, "null renderer");    \
    paths_status st = (expr);                                       \
    return st == PATHS_OK ? st : fail(st, r->impl->err);            \
    PATHS_GUARD_END

paths_status paths_renderer_render(paths_renderer *r, uint32_t spp) { PATHS_FORWARD(r->impl->render(spp, true)) }
paths_status paths_renderer_render_async(paths_renderer *r, uint32_t spp) { PATHS_FORWARD(r->impl->render(spp, false)) }
paths_status paths_renderer_sync(paths_renderer *r) { PATHS_FORWARD(r->impl->sync()) }
paths_status paths_renderer_combine(paths_renderer *r, int32_t root) { PATHS_FORWARD(r->impl->combine(root, true)) }
paths_status paths_renderer_combine_async(paths_renderer *r, int32_t root) { PATHS_FORWARD(r->impl->combine(root, false)) }
paths_status paths_renderer_image_f32(paths_renderer *r, float *rgb) {
    if (!rgb) return fail(PATHS_ERR_INVALID_ARG, "null buffer");
    PATHS_FORWARD(r->impl->image_f32(rgb))
}
paths_status paths_renderer_image_rgb24(paths_renderer *r, uint8_t *rgb) {
    if (!rgb) return fail(PATHS_ERR_INVALID_ARG, "null buffer");
    PATHS_FORWARD(r->impl->image_rgb24(rgb))
}
paths_status paths_renderer_get_sums(paths_renderer *r, double *sums4) {
    if (!sums4) return fail(PATHS_ERR_INVALID_ARG, "null buffer");
    PATHS_FORWARD(r->impl->get_sums(sums4))
}
paths_status paths_renderer_add_sums(paths_renderer *r, const double *sums4) {
    if (!sums4) return fail(PATHS_ERR_INVALID_ARG, "null buffer");
    PATHS_FORWARD(r->impl->add_sums(sums4))
}
paths_status paths_renderer_device_sums(paths_renderer *r, void **device_ptr, size_t *n_elems, int32_t *elem_bytes) {
    if (!device_ptr || !n_elems || !elem_bytes) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    PATHS_FORWARD(r->impl->device_sums(device_ptr, n_elems, elem_bytes))
}
paths_status paths_renderer_cuda_stream(paths_renderer *r, void **stream) {
    if (!stream) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    PATHS_FORWARD(r->impl->cuda_stream(stream))
}
paths_status paths_renderer_get_stats(paths_renderer *r, paths_stats *st_out) {
    if (!st_out) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    PATHS_FORWARD(r->impl->get_stats(st_out))
}
paths_status paths_renderer_set_camera(paths_renderer *r, const double location[3], const double rot[9]) {
    if (!location || !rot) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    PATHS_FORWARD(r->impl->set_camera(location, rot))
}
paths_status paths_renderer_reset(paths_renderer *r) { PATHS_FORWARD(r->impl->reset()) }
paths_status paths_renderer_render_preview(paths_renderer *r) { PATHS_FORWARD(r->impl->render_preview()) }
paths_status paths_renderer_render_columns(paths_renderer *r, uint32_t n_requests) { PATHS_FORWARD(r->impl->render_columns(n_requests)) }
paths_status paths_renderer_get_progress(paths_renderer *r, uint32_t *samples_per_pixel, uint32_t *sweep_position) {
    if (!samples_per_pixel || !sweep_position) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    PATHS_FORWARD(r->impl->get_progress(samples_per_pixel, sweep_position))
}

// ---- checkpoint / resume: the estimator (sum + count per pixel) and the position in the sample schedule are the whole state
namespace {
struct CheckpointHeader {
    char magic[8]; // "PATHSCK2"
    uint32_t width, height;
    uint32_t samples_done;
    uint32_t max_bounces;
    uint64_t seed;
    // everything else that decides which estimator the sums belong to: resuming under other settings would silently mix estimators
    uint32_t extensions;
    int32_t precision, shard_mode;
    uint32_t rank, world;
    uint32_t reserved;
};
} // namespace

paths_status paths_renderer_save_checkpoint(paths_renderer *r, const char *path) {
    PATHS_GUARD_BEGIN
    if (!r || !path) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    uint32_t done = 0, sweep = 0;
    paths_status st = r->impl->get_progress(&done, &sweep);
    if (st != PATHS_OK) return fail(st, r->impl->err);
    if (sweep != 0) return fail(PATHS_ERR_INVALID_ARG, "a column sweep is in progress: checkpoints are taken between sweeps");
    std::vector<double> sums((size_t)r->w * r->h * 4);
    st = r->impl->get_sums(sums.data());
    if (st != PATHS_OK) return fail(st, r->impl->err);
    CheckpointHeader h;
    std::memset(&h, 0, sizeof(h));
    std::memcpy(h.magic, "PATHSCK2", 8);
    h.width = r->w;
    h.height = r->h;
    h.samples_done = done;
    h.seed = r->seed;
    h.max_bounces = r->opts.max_bounces;
    h.extensions = r->opts.extensions;
    h.precision = r->opts.precision;
    h.shard_mode = r->world > 1 ? r->opts.shard_mode : 0;
    h.rank = r->rank;
    h.world = r->world;
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(PATHS_ERR_IO, std::string("cannot open ") + path);
    bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1 && std::fwrite(sums.data(), sizeof(double), sums.size(), f) == sums.size();
    ok = (std::fclose(f) == 0) && ok;
    return ok ? PATHS_OK : fail(PATHS_ERR_IO, std::string("short write to ") + path);
    PATHS_GUARD_END
}

paths_status paths_renderer_load_checkpoint(paths_renderer *r, const char *path) {
    PATHS_GUARD_BEGIN
    if (!r || !path) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    FILE *f = std::fopen(path, "rb");
    if (!f) return fail(PATHS_ERR_IO, std::string("cannot open ") + path);
    CheckpointHeader h;
    std::vector<double> sums;
    bool ok = std::fread(&h, sizeof(h), 1, f) == 1 && std::memcmp(h.magic, "PATHSCK2", 8) == 0;
    if (ok && (h.width != r->w || h.height != r->h)) {
        std::fclose(f);
        return fail(PATHS_ERR_INVALID_ARG, "checkpoint was taken at another image size");
    }
    if (ok && h.seed != r->seed) {
        std::fclose(f);
        return fail(PATHS_ERR_INVALID_ARG, "checkpoint was taken with another seed: resuming would repeat or skip samples");
    }
    if (ok && (h.max_bounces != r->opts.max_bounces || h.extensions != r->opts.extensions || h.precision != r->opts.precision ||
               h.world != r->world || h.rank != r->rank || (r->world > 1 && h.shard_mode != r->opts.shard_mode))) {
        std::fclose(f);
        return fail(PATHS_ERR_INVALID_ARG, "checkpoint was taken under other render options (max_bounces, extensions, precision or shard layout): "
                                           "resuming would mix two estimators");
    }
    if (ok) {
        sums.resize((size_t)h.width * h.height * 4);
        ok = std::fread(sums.data(), sizeof(double), sums.size(), f) == sums.size();
    }
    std::fclose(f);
    if (!ok) return fail(PATHS_ERR_PARSE, std::string("not a checkpoint file: ") + path);
    paths_status st = r->impl->reset(); // new epoch, then the saved estimator on top of zeros
    if (st == PATHS_OK) st = r->impl->add_sums(sums.data());
    if (st == PATHS_OK) st = r->impl->set_progress(h.samples_done);
    return st == PATHS_OK ? st : fail(st, r->impl->err);
    PATHS_GUARD_END
}
void paths_renderer_destroy(paths_renderer *r) { delete r; }

paths_status paths_camera_rays(const paths_camera *cam, int32_t device, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0, uint32_t w,
                               uint32_t h, double *o, double *d, double *weight, int32_t precision) {
    PATHS_GUARD_BEGIN
    if (!cam || !o || !d || !weight) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    if ((uint64_t)x0 + w > cam->image_width || (uint64_t)y0 + h > cam->image_height) return fail(PATHS_ERR_INVALID_ARG, "rectangle outside the image");
    paths_status st = check_device(device);
    if (st != PATHS_OK) return st;
    std::string err;
    st = precision == PATHS_PRECISION_F32 ? camera_rays_f32(*cam, device, seed, sample, x0, y0, w, h, o, d, weight, err)
                                          : camera_rays_f64(*cam, device, seed, sample, x0, y0, w, h, o, d, weight, err);
    return st == PATHS_OK ? st : fail(st, err);
    PATHS_GUARD_END
}

paths_status paths_write_ppm(const char *path, const uint8_t *rgb, uint32_t w, uint32_t h) {
    if (!path || !rgb) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(PATHS_ERR_IO, std::string("cannot open ") + path);
    std::fprintf(f, "P6\n%u %u\n255\n", w, h);
    size_t n = (size_t)w * h * 3;
    bool ok = std::fwrite(rgb, 1, n, f) == n;
    ok = (std::fclose(f) == 0) && ok;
    return ok ? PATHS_OK : fail(PATHS_ERR_IO, std::string("short write to ") + path);
}

paths_status paths_write_pfm(const char *path, const float *rgb, uint32_t w, uint32_t h) {
    if (!path || !rgb) return fail(PATHS_ERR_INVALID_ARG, "null argument");
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(PATHS_ERR_IO, std::string("cannot open ") + path);
    std::fprintf(f, "PF\n%u %u\n-1.0\n", w, h);
    bool ok = true;
    for (uint32_t y = 0; y < h && ok; y++) // PFM stores the bottom row first
        ok = std::fwrite(rgb + (size_t)(h - 1 - y) * w * 3, sizeof(float), (size_t)w * 3, f) == (size_t)w * 3;
    ok = (std::fclose(f) == 0) && ok;
    return ok ? PATHS_OK : fail(PATHS_ERR_IO, std::string("short write to ") + path);
}

} // extern "C"
