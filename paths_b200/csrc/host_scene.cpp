// host_scene.cpp -- SceneDescription::scene + Scene::new on the host (src/serde.rs:81-154, src/scene.rs:143-170,
// src/model.rs:174-249, src/geom.rs:98-122,198-205,251-261).  f64 throughout; build with -ffp-contract=off.
#include "host_scene.h"

#include <cmath>
#include <cstring>
#include <limits>

namespace paths {

D3 normed(D3 a) { return a / std::sqrt(dot(a, a)); } // src/vector.rs:39-41 (magnitude() is the squared length)
bool has_nan(D3 a) { return std::isnan(a.x) || std::isnan(a.y) || std::isnan(a.z); }

static Mat3 mat_mul(const Mat3 &a, const Mat3 &b) { // src/matrix.rs:78-92
    Mat3 o;
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
            double acc = 0.0;
            for (int k = 0; k < 3; k++) acc += a.m[r * 3 + k] * b.m[k * 3 + c];
            o.m[r * 3 + c] = acc;
        }
    return o;
}

Mat3 mat_rotation(double yaw, double pitch, double roll) {
    double sp = std::sin(pitch), cp = std::cos(pitch);
    double sy = std::sin(yaw), cy = std::cos(yaw);
    double sr = std::sin(roll), cr = std::cos(roll);
    Mat3 rx{{1.0, 0.0, 0.0, 0.0, cp, -sp, 0.0, sp, cp}};
    Mat3 ry{{cy, 0.0, sy, 0.0, 1.0, 0.0, -sy, 0.0, cy}};
    Mat3 rz{{cr, -sr, 0.0, sr, cr, 0.0, 0.0, 0.0, 1.0}};
    return mat_mul(mat_mul(rx, ry), rz);
}

D3 mat_apply(const Mat3 &m, D3 v) {
    return d3(m.m[0] * v.x + m.m[1] * v.y + m.m[2] * v.z, m.m[3] * v.x + m.m[4] * v.y + m.m[5] * v.z,
              m.m[6] * v.x + m.m[7] * v.y + m.m[8] * v.z);
}

static void fill_basic(const paths_basic_material &b, int &type, double albedo[3], double &r0, double &metal, double &rough) {
    type = b.type;
    for (int i = 0; i < 3; i++) albedo[i] = b.albedo[i];
    r0 = b.reflectance;
    metal = b.metalness;
    rough = b.roughness;
}

static bool convert_material(const paths_material_desc &d, HostMaterial &m, std::string &err) {
    std::memset(&m, 0, sizeof(m));
    if (d.m.type < PATHS_MAT_LAMBERTIAN || d.m.type > PATHS_MAT_FRESNEL) {
        err = "unknown material type";
        return false;
    }
    fill_basic(d.m, m.type, m.albedo, m.r0, m.metalness, m.roughness);
    m.vertex_albedo = d.m.albedo_mode == 1;
    if (d.m.type == PATHS_MAT_FRESNEL) {
        fill_basic(d.diffuse, m.d_type, m.d_albedo, m.d_r0, m.d_metal, m.d_rough);
        fill_basic(d.specular, m.s_type, m.s_albedo, m.s_r0, m.s_metal, m.s_rough);
        double n1 = 1.0, n2 = d.refractive_index; // src/material.rs:380-387
        m.fresnel_r0 = std::pow((n1 - n2) / (n1 + n2), 2.0);
        m.vertex_albedo = 0;
    }
    return true;
}

static Box3 sphere_box(D3 c, double r) { return Box3{c - d3(r, r, r), c + d3(r, r, r)}; }
static Box3 tri_box(D3 a, D3 b, D3 c) {
    Box3 o;
    o.lo = d3(std::fmin(a.x, std::fmin(b.x, c.x)), std::fmin(a.y, std::fmin(b.y, c.y)), std::fmin(a.z, std::fmin(b.z, c.z)));
    o.hi = d3(std::fmax(a.x, std::fmax(b.x, c.x)), std::fmax(a.y, std::fmax(b.y, c.y)), std::fmax(a.z, std::fmax(b.z, c.z)));
    return o;
}

bool build_host_scene(const paths_scene_desc &desc, HostScene &out, std::string &err) {
    out = HostScene();
    // ---- models: face normals now, vertex normals on demand (src/model.rs:116-128)
    std::vector<std::vector<D3>> face_normals(desc.n_models);
    std::vector<std::vector<D3>> verts(desc.n_models);
    for (uint32_t mi = 0; mi < desc.n_models; mi++) {
        const paths_model_desc &md = desc.models[mi];
        if ((md.n_vertices && !md.vertices) || (md.n_faces && !md.faces)) {
            err = "model with null arrays";
            return false;
        }
        HostModel hm;
        hm.vert_base = (uint32_t)out.vertex_normals.size();
        hm.face_base = (uint32_t)(out.faces.size() / 3);
        hm.n_verts = md.n_vertices;
        hm.n_faces = md.n_faces;
        hm.has_normals = false;
        hm.has_colours = md.vertex_colours != nullptr;
        verts[mi].resize(md.n_vertices);
        for (uint32_t v = 0; v < md.n_vertices; v++) verts[mi][v] = d3(md.vertices[3 * v], md.vertices[3 * v + 1], md.vertices[3 * v + 2]);
        face_normals[mi].resize(md.n_faces);
        for (uint32_t f = 0; f < md.n_faces; f++) {
            uint32_t ia = md.faces[3 * f], ib = md.faces[3 * f + 1], ic = md.faces[3 * f + 2];
            if (ia >= md.n_vertices || ib >= md.n_vertices || ic >= md.n_vertices) {
                err = "face index out of range";
                return false;
            }
            out.faces.push_back(ia);
            out.faces.push_back(ib);
            out.faces.push_back(ic);
            D3 v1 = verts[mi][ia], v2 = verts[mi][ib], v3 = verts[mi][ic];
            D3 s1 = v2 - v1, s2 = v3 - v1, s3 = v3 - v2; // src/model.rs:237-249
            D3 n = normed(cross(s1, s2));
            if (has_nan(n)) n = normed(cross(s1, s3));
            face_normals[mi][f] = n;
        }
        const double qnan = std::numeric_limits<double>::quiet_NaN();
        out.vertex_normals.resize(out.vertex_normals.size() + md.n_vertices, d3(qnan, qnan, qnan));
        out.vertex_colours.resize(out.vertex_colours.size() + (size_t)3 * md.n_vertices, 0.0);
        if (md.vertex_colours)
            for (size_t k = 0; k < (size_t)3 * md.n_vertices; k++) out.vertex_colours[(size_t)3 * hm.vert_base + k] = md.vertex_colours[k];
        out.models.push_back(hm);
    }
    // ---- objects in order; mesh faces in face order (src/scene.rs:144-154)
    for (uint32_t oi = 0; oi < desc.n_objects; oi++) {
        const paths_object_desc &od = desc.objects[oi];
        HostObject ho;
        ho.shape = od.shape;
        ho.model = od.model;
        ho.smooth = od.smooth_normals != 0;
        // src/serde.rs:103 passes (pitch, yaw, roll) to rotation(yaw, pitch, roll): the YAML pitch drives Ry, the YAML yaw Rx.
        ho.rot = mat_rotation(od.rot_pitch, od.rot_yaw, od.rot_roll);
        if (!convert_material(od.material, ho.mat, err)) return false;
        if (od.shape == 0) {
            HostPrim p;
            std::memset(&p, 0, sizeof(p));
            p.type = 0;
            p.entity = oi;
            p.a = d3(od.center[0], od.center[1], od.center[2]);
            p.b = d3(od.radius, 0.0, 0.0);
            p.box = sphere_box(p.a, od.radius);
            out.prims.push_back(p);
            ho.smooth = 0;
        } else if (od.shape == 1) {
            if (od.model < 0 || (uint32_t)od.model >= desc.n_models) {
                err = "object model index out of range";
                return false;
            }
            HostModel &hm = out.models[od.model];
            const std::vector<D3> &fn = face_normals[od.model];
            const std::vector<D3> &vv = verts[od.model];
            if (ho.mat.vertex_albedo && !hm.has_colours) {
                err = "Model does not have vertex colours"; // src/model.rs:170
                return false;
            }
            if (ho.smooth && !hm.has_normals) { // src/serde.rs:115-120 -> src/model.rs:194-224
                std::vector<D3> sums(hm.n_verts, d3(0.0, 0.0, 0.0));
                std::vector<uint64_t> counts(hm.n_verts, 0);
                for (uint32_t f = 0; f < hm.n_faces; f++) {
                    D3 n = fn[f];
                    if (has_nan(n)) continue;
                    for (int k = 0; k < 3; k++) {
                        uint32_t v = out.faces[3 * ((size_t)hm.face_base + f) + k];
                        sums[v] = sums[v] + n;
                        counts[v] += 1;
                    }
                }
                for (uint32_t v = 0; v < hm.n_verts; v++) out.vertex_normals[hm.vert_base + v] = sums[v] / (double)counts[v];
                hm.has_normals = true;
            }
            D3 tr = d3(od.translation[0], od.translation[1], od.translation[2]);
            for (uint32_t f = 0; f < hm.n_faces; f++) {
                if (has_nan(fn[f])) continue; // src/model.rs:185-189
                HostPrim p;
                std::memset(&p, 0, sizeof(p));
                p.type = 1;
                p.entity = oi;
                p.face = f;
                const uint32_t *fi = &out.faces[3 * ((size_t)hm.face_base + f)];
                // src/geom.rs:251-261: vertex' = (R v) * scale + t ; normal' = R n
                p.a = mat_apply(ho.rot, vv[fi[0]]) * od.scale + tr;
                p.b = mat_apply(ho.rot, vv[fi[1]]) * od.scale + tr;
                p.c = mat_apply(ho.rot, vv[fi[2]]) * od.scale + tr;
                p.n = mat_apply(ho.rot, fn[f]);
                p.box = tri_box(p.a, p.b, p.c);
                out.prims.push_back(p);
            }
        } else {
            err = "unknown shape";
            return false;
        }
        out.objects.push_back(ho);
    }
    // ---- lights; area lights append their sphere (src/scene.rs:155-164)
    for (uint32_t li = 0; li < desc.n_lights; li++) {
        const paths_light_desc &ld = desc.lights[li];
        HostLight hl;
        hl.geometry = ld.geometry;
        hl.v = d3(ld.v[0], ld.v[1], ld.v[2]);
        hl.radius = ld.radius;
        for (int i = 0; i < 3; i++) hl.colour[i] = ld.colour[i];
        hl.intensity = ld.intensity;
        hl.prim = PATHS_MISS;
        if (ld.geometry == 1) {
            HostPrim p;
            std::memset(&p, 0, sizeof(p));
            p.type = 0;
            p.is_light = 1;
            p.entity = li;
            p.a = hl.v;
            p.b = d3(ld.radius, 0.0, 0.0);
            p.box = sphere_box(p.a, ld.radius);
            hl.prim = (uint32_t)out.prims.size();
            out.prims.push_back(p);
        } else if (ld.geometry != 0) {
            err = "unknown light geometry";
            return false;
        }
        out.lights.push_back(hl);
    }
    if (out.prims.empty()) {
        err = "Must have at least one cluster (scene has no primitives)"; // src/bvh.rs:201
        return false;
    }
    out.sky_type = desc.sky.type;
    for (int i = 0; i < 3; i++) {
        out.sky_c0[i] = desc.sky.c0[i];
        out.sky_c1[i] = desc.sky.c1[i];
    }
    if (desc.sky.type == 2) {
        if (!desc.sky.texels || !desc.sky.width || !desc.sky.height) {
            err = "Hdri sky without texels";
            return false;
        }
        out.sky_w = desc.sky.width;
        out.sky_h = desc.sky.height;
        out.sky_texels.assign(desc.sky.texels, desc.sky.texels + (size_t)3 * desc.sky.width * desc.sky.height);
        // piecewise-constant importance function: texel luminance x solid angle (row y covers latitude pi (1 - (y + .5) / h),
        // the mapping of src/scene.rs:95-111)
        const uint32_t w = out.sky_w, h = out.sky_h;
        out.env_marginal.assign((size_t)h + 1, 0.0f);
        out.env_conditional.assign((size_t)h * (w + 1), 0.0f);
        std::vector<double> row_sum(h, 0.0);
        const double pi = 3.14159265358979323846;
        for (uint32_t y = 0; y < h; y++) {
            double sin_lat = std::sin(pi * (1.0 - (y + 0.5) / h));
            double acc = 0.0;
            float *cond = &out.env_conditional[(size_t)y * (w + 1)];
            for (uint32_t x = 0; x < w; x++) {
                const float *t = &out.sky_texels[3 * ((size_t)y * w + x)];
                double lum = std::fmax(0.0, ((double)t[0] + t[1] + t[2]) / 3.0) + 1e-12;
                acc += lum;
                cond[x + 1] = (float)acc;
            }
            for (uint32_t x = 1; x <= w; x++) cond[x] = (float)(cond[x] / acc);
            cond[w] = 1.0f;
            row_sum[y] = acc * sin_lat;
        }
        double tot = 0.0;
        for (uint32_t y = 0; y < h; y++) {
            tot += row_sum[y];
            out.env_marginal[y + 1] = (float)tot;
        }
        for (uint32_t y = 1; y <= h; y++) out.env_marginal[y] = (float)(out.env_marginal[y] / tot);
        out.env_marginal[h] = 1.0f;
    } else if (desc.sky.type != 0 && desc.sky.type != 1) {
        err = "unknown skybox type";
        return false;
    }
    // ---- statistics; which precision the scene needs
    const double inf = std::numeric_limits<double>::infinity();
    out.bounds_lo = d3(inf, inf, inf);
    out.bounds_hi = d3(-inf, -inf, -inf);
    for (const HostPrim &p : out.prims) {
        if (p.type == 0) out.n_spheres++;
        else out.n_tris++;
        if (p.type == 1 && (p.box.lo.x == p.box.hi.x || p.box.lo.y == p.box.hi.y || p.box.lo.z == p.box.hi.z)) out.hidden++;
        if (p.type == 0 && p.b.x >= 1e5) continue; // "plane" spheres
        out.bounds_lo = d3(std::fmin(out.bounds_lo.x, p.box.lo.x), std::fmin(out.bounds_lo.y, p.box.lo.y), std::fmin(out.bounds_lo.z, p.box.lo.z));
        out.bounds_hi = d3(std::fmax(out.bounds_hi.x, p.box.hi.x), std::fmax(out.bounds_hi.y, p.box.hi.y), std::fmax(out.bounds_hi.z, p.box.hi.z));
    }
    double extent = 0.0;
    if (out.bounds_lo.x <= out.bounds_hi.x)
        extent = std::fmax(std::fmax(std::fabs(out.bounds_lo.x), std::fabs(out.bounds_hi.x)),
                           std::fmax(std::fmax(std::fabs(out.bounds_lo.y), std::fabs(out.bounds_hi.y)),
                                     std::fmax(std::fabs(out.bounds_lo.z), std::fabs(out.bounds_hi.z))));
    // The integrator offsets new ray origins by an absolute 1e-4 (src/trace.rs:57,89).  f32 origins keep that offset
    // meaningful while ulp(extent) is an order of magnitude below it: ulp(128) = 1.5e-5.
    out.auto_precision = extent > 128.0 ? PATHS_PRECISION_MIXED : PATHS_PRECISION_F32;
    return true;
}

} // namespace paths
