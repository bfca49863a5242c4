// scene_device.cu -- stages a HostScene into HBM (layout: DESIGN.md "Data layout in HBM").
#include <cmath>
#include <cstring>

#include "scene_device.h"

namespace paths {

SceneDevice::~SceneDevice() {
    if (device >= 0) cudaSetDevice(device);
    for (void *p : allocs) cudaFree(p);
}

namespace {
template <typename T> bool put(SceneDevice *sd, const std::vector<T> &src, const T **dst, std::string &err) {
    *dst = nullptr;
    size_t bytes = std::max<size_t>(src.size(), 1) * sizeof(T);
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
        err = std::string("cudaMalloc: ") + cudaGetErrorString(e);
        return false;
    }
    sd->allocs.push_back(p);
    sd->bytes += bytes;
    if (!src.empty()) {
        e = cudaMemcpy(p, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            err = std::string("cudaMemcpy: ") + cudaGetErrorString(e);
            return false;
        }
    }
    *dst = (const T *)p;
    return true;
}
} // namespace

DevCamera make_dev_camera(const paths_camera &c) {
    DevCamera d;
    d.width = c.image_width;
    d.height = c.image_height;
    for (int i = 0; i < 3; i++) d.location[i] = c.location[i];
    Mat3 r = mat_rotation(c.yaw, c.pitch, c.roll); // serde.rs:177
    for (int i = 0; i < 9; i++) d.rot[i] = r.m[i];
    d.sensor_width = c.sensor_width;
    d.sensor_height = c.sensor_height;
    d.focal_length = c.focal_length;
    d.aperture = c.aperture;
    d.distance_from_lens = (c.focal_length * c.focus_distance) / (c.focus_distance - c.focal_length); // serde.rs:185
    return d;
}

SceneDevice *upload_scene(int device, const HostScene &hs, const RefTree *ref, const FastTree *fast, std::string &err, bool want_f32_tables) {
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        err = std::string("cudaSetDevice: ") + cudaGetErrorString(e);
        return nullptr;
    }
    SceneDevice *sd = new SceneDevice();
    sd->device = device;
    DevScene &v = sd->view;
    std::memset(&v, 0, sizeof(v));
    v.n_prims = (uint32_t)hs.prims.size();
    v.n_lights = (uint32_t)hs.lights.size();
    v.n_objects = (uint32_t)hs.objects.size();
    bool ok = true;
    { // prim tables
        std::vector<uint4> meta(hs.prims.size());
        std::vector<double> geom(hs.prims.size() * 12, 0.0);
        for (size_t i = 0; i < hs.prims.size(); i++) {
            const HostPrim &p = hs.prims[i];
            meta[i] = make_uint4((p.type == 1 ? PATHS_PM_TRI : 0u) | (p.is_light ? PATHS_PM_LIGHT : 0u), p.entity, p.face, 0u);
            double *g = &geom[i * 12];
            if (p.type == 1) {
                g[0] = p.a.x; g[1] = p.a.y; g[2] = p.a.z; g[3] = p.b.x; g[4] = p.b.y; g[5] = p.b.z;
                g[6] = p.c.x; g[7] = p.c.y; g[8] = p.c.z; g[9] = p.n.x; g[10] = p.n.y; g[11] = p.n.z;
            } else {
                g[0] = p.a.x; g[1] = p.a.y; g[2] = p.a.z; g[3] = p.b.x;
            }
        }
        ok = ok && put(sd, meta, &v.prim_meta, err) && put(sd, geom, &v.prim_geom, err);
    }
    if (ok && ref) {
        std::vector<DevRefNode> nodes(ref->nodes.size());
        for (size_t i = 0; i < nodes.size(); i++) {
            std::memcpy(nodes[i].lbox, ref->nodes[i].lbox, sizeof(double) * 6);
            std::memcpy(nodes[i].rbox, ref->nodes[i].rbox, sizeof(double) * 6);
            nodes[i].left = ref->nodes[i].left;
            nodes[i].right = ref->nodes[i].right;
            nodes[i].pad[0] = nodes[i].pad[1] = 0;
        }
        ok = ok && put(sd, nodes, &v.ref_nodes, err);
        v.ref_root = ref->root;
        std::memcpy(v.ref_root_box, ref->root_box, sizeof(double) * 6);
    }
    if (ok && fast) {
        std::vector<float4> nodes(fast->nodes.size() * 4);
        for (size_t i = 0; i < fast->nodes.size(); i++) {
            const FastNode2 &n = fast->nodes[i];
            nodes[4 * i + 0] = make_float4(n.c0[0], n.c0[1], n.c0[2], n.c0[3]);
            nodes[4 * i + 1] = make_float4(n.c1[0], n.c1[1], n.c1[2], n.c1[3]);
            nodes[4 * i + 2] = make_float4(n.c0[4], n.c0[5], n.c1[4], n.c1[5]);
            float4 c;
            std::memcpy(&c.x, &n.i0, 4);
            std::memcpy(&c.y, &n.i1, 4);
            c.z = c.w = 0.0f;
            nodes[4 * i + 3] = c;
        }
        std::vector<float4> leaves(fast->leaf_prims.size() * 3);
        for (size_t i = 0; i < fast->leaf_prims.size(); i++) {
            uint32_t prim = fast->leaf_prims[i];
            const HostPrim &p = hs.prims[prim];
            float4 *s = &leaves[3 * i];
            if (p.type == 1) {
                uint32_t tag = prim;
                s[0] = make_float4((float)p.a.x, (float)p.a.y, (float)p.a.z, 0.0f);
                s[1] = make_float4((float)p.b.x, (float)p.b.y, (float)p.b.z, 0.0f);
                s[2] = make_float4((float)p.c.x, (float)p.c.y, (float)p.c.z, 0.0f);
                std::memcpy(&s[2].w, &tag, 4);
            } else {
                double c[4] = {p.a.x, p.a.y, p.a.z, p.b.x};
                std::memcpy(&s[0], c, 32);
                uint32_t tag = prim | 0x80000000u;
                s[2] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                std::memcpy(&s[2].w, &tag, 4);
            }
        }
        ok = ok && put(sd, nodes, &v.fast_nodes, err) && put(sd, leaves, &v.fast_leaves, err);
        v.fast_root = fast->root;
        static_assert(sizeof(FastNode4) == 128, "FastNode4 must be one cache line");
        std::vector<float4> nodes4(fast->nodes4.size() * 8);
        if (!fast->nodes4.empty()) std::memcpy(nodes4.data(), fast->nodes4.data(), fast->nodes4.size() * sizeof(FastNode4));
        ok = ok && put(sd, nodes4, &v.fast4_nodes, err);
        v.fast4_root = fast->root4;
        v.fast4_depth = fast->depth4;
    }
    if (ok) { // shading tables
        std::vector<DevObject> objs(hs.objects.size());
        for (size_t i = 0; i < objs.size(); i++) {
            const HostObject &o = hs.objects[i];
            DevObject &d = objs[i];
            std::memset(&d, 0, sizeof(d));
            d.shape = o.shape;
            d.smooth = o.smooth;
            if (o.shape == 1) {
                d.vert_base = hs.models[o.model].vert_base;
                d.face_base = hs.models[o.model].face_base;
            }
            for (int k = 0; k < 9; k++) d.rot[k] = o.rot.m[k];
            const HostMaterial &m = o.mat;
            d.mat_type = m.type;
            d.vertex_albedo = m.vertex_albedo;
            for (int k = 0; k < 3; k++) {
                d.albedo[k] = m.albedo[k];
                d.d_albedo[k] = m.d_albedo[k];
                d.s_albedo[k] = m.s_albedo[k];
            }
            d.r0 = m.r0; d.metalness = m.metalness; d.roughness = m.roughness;
            d.d_type = m.d_type; d.s_type = m.s_type;
            d.d_r0 = m.d_r0; d.d_metal = m.d_metal; d.d_rough = m.d_rough;
            d.s_r0 = m.s_r0; d.s_metal = m.s_metal; d.s_rough = m.s_rough;
            d.fresnel_r0 = m.fresnel_r0;
        }
        std::vector<DevLight> lights(hs.lights.size());
        for (size_t i = 0; i < lights.size(); i++) {
            const HostLight &l = hs.lights[i];
            DevLight &d = lights[i];
            d.geometry = l.geometry;
            d.prim = l.prim;
            d.v[0] = l.v.x; d.v[1] = l.v.y; d.v[2] = l.v.z;
            d.radius = l.radius;
            for (int k = 0; k < 3; k++) d.colour[k] = l.colour[k];
            d.intensity = l.intensity;
        }
        std::vector<double> vn(hs.vertex_normals.size() * 3);
        for (size_t i = 0; i < hs.vertex_normals.size(); i++) {
            vn[3 * i] = hs.vertex_normals[i].x;
            vn[3 * i + 1] = hs.vertex_normals[i].y;
            vn[3 * i + 2] = hs.vertex_normals[i].z;
        }
        if (fast || want_f32_tables) { // f32 mirrors for the production shade kernel
            std::vector<DevObjectF> objs32(objs.size());
            for (size_t i = 0; i < objs.size(); i++) {
                const DevObject &a = objs[i];
                DevObjectF &b = objs32[i];
                std::memset(&b, 0, sizeof(b));
                b.shape = a.shape; b.smooth = a.smooth; b.vert_base = a.vert_base; b.face_base = a.face_base;
                for (int k = 0; k < 9; k++) b.rot[k] = (float)a.rot[k];
                b.mat_type = a.mat_type; b.vertex_albedo = a.vertex_albedo;
                for (int k = 0; k < 3; k++) {
                    b.albedo[k] = (float)a.albedo[k];
                    b.d_albedo[k] = (float)a.d_albedo[k];
                    b.s_albedo[k] = (float)a.s_albedo[k];
                }
                b.r0 = (float)a.r0; b.metalness = (float)a.metalness; b.roughness = (float)a.roughness;
                b.d_type = a.d_type; b.s_type = a.s_type;
                b.d_r0 = (float)a.d_r0; b.d_metal = (float)a.d_metal; b.d_rough = (float)a.d_rough;
                b.s_r0 = (float)a.s_r0; b.s_metal = (float)a.s_metal; b.s_rough = (float)a.s_rough;
                b.fresnel_r0 = (float)a.fresnel_r0;
            }
            std::vector<float4> vn32(hs.vertex_normals.size()), vc32(hs.vertex_colours.size() / 3), ps32(hs.prims.size() * 2);
            for (size_t i = 0; i < vn32.size(); i++)
                vn32[i] = make_float4((float)hs.vertex_normals[i].x, (float)hs.vertex_normals[i].y, (float)hs.vertex_normals[i].z, 0.0f);
            for (size_t i = 0; i < vc32.size(); i++)
                vc32[i] = make_float4((float)hs.vertex_colours[3 * i], (float)hs.vertex_colours[3 * i + 1], (float)hs.vertex_colours[3 * i + 2], 0.0f);
            for (size_t i = 0; i < hs.prims.size(); i++) {
                const HostPrim &p = hs.prims[i];
                uint32_t m[4] = {(p.type == 1 ? PATHS_PM_TRI : 0u) | (p.is_light ? PATHS_PM_LIGHT : 0u), p.entity, p.face, 0u};
                std::memcpy(&ps32[2 * i], m, 16);
                const D3 &g = p.type == 1 ? p.n : p.a;
                ps32[2 * i + 1] = make_float4((float)g.x, (float)g.y, (float)g.z, 0.0f);
            }
            ok = ok && put(sd, objs32, &v.objects32, err) && put(sd, vn32, &v.vertex_normals32, err) && put(sd, vc32, &v.vertex_colours32, err) &&
                 put(sd, ps32, &v.prim_shade, err);
        }
        ok = ok && put(sd, objs, &v.objects, err) && put(sd, lights, &v.lights, err) && put(sd, vn, &v.vertex_normals, err) &&
             put(sd, hs.vertex_colours, &v.vertex_colours, err) && put(sd, hs.faces, &v.faces, err) && put(sd, hs.sky_texels, &v.sky_texels, err) &&
             put(sd, hs.env_marginal, &v.env_marginal, err) && put(sd, hs.env_conditional, &v.env_conditional, err);
        v.sky_type = hs.sky_type;
        for (int k = 0; k < 3; k++) {
            v.sky_c0[k] = hs.sky_c0[k];
            v.sky_c1[k] = hs.sky_c1[k];
        }
        v.sky_w = hs.sky_w;
        v.sky_h = hs.sky_h;
    }
    if (!ok) {
        delete sd;
        return nullptr;
    }
    return sd;
}

} // namespace paths
