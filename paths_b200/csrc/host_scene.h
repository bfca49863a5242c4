// host_scene.h -- host-side scene model of the B200 render path: the flattened primitive list in the reference's item
// order (src/scene.rs:143-170), shading tables, and the two acceleration structures that are staged into HBM.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/paths_b200.h"

namespace paths {

struct D3 {
    double x, y, z;
};
inline D3 d3(double x, double y, double z) { return D3{x, y, z}; }
inline D3 operator+(D3 a, D3 b) { return d3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline D3 operator-(D3 a, D3 b) { return d3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline D3 operator*(D3 a, double s) { return d3(a.x * s, a.y * s, a.z * s); }
inline D3 operator/(D3 a, double s) { return d3(a.x / s, a.y / s, a.z / s); }
inline double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// src/vector.rs:43-49 (the y component is written -(x*oz - z*ox); same value, same rounding)
inline D3 cross(D3 a, D3 b) { return d3(a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x); }
D3 normed(D3 a);
bool has_nan(D3 a);

struct Mat3 {
    double m[9]; // row-major
};
Mat3 mat_rotation(double yaw, double pitch, double roll); // src/matrix.rs:30-35
D3 mat_apply(const Mat3 &m, D3 v);                        // src/matrix.rs:94-104

struct Box3 {
    D3 lo, hi;
};

// one flattened item of Scene::new (src/scene.rs:143-170): prim id == index == leaf.obj of the reference BVH
struct HostPrim {
    int type;        // 0 sphere, 1 triangle
    int is_light;    // EntityID::Light / EntityID::Object
    uint32_t entity; // object id or light id
    uint32_t face;   // triangle: face index inside its model (CollisionMetadata, src/geom.rs:48-57)
    D3 a, b, c;      // triangle vertices (world); sphere: a = centre, b.x = radius
    D3 n;            // triangle: rotated face normal
    Box3 box;        // src/geom.rs:237-240, 305-320
};

struct HostMaterial {
    int type; // PATHS_MAT_*; FRESNEL and COOK_TORRANCE cannot be sampled (src/material.rs:81-88)
    int vertex_albedo;
    double albedo[3];
    double r0, metalness, roughness;
    // Fresnel combination (src/material.rs:373-421): brdf() only
    int d_type, s_type;
    double d_albedo[3], s_albedo[3];
    double d_r0, d_metal, d_rough, s_r0, s_metal, s_rough;
    double fresnel_r0;
};

struct HostObject {
    int shape; // 0 sphere, 1 mesh
    int model;
    int smooth;
    Mat3 rot;
    HostMaterial mat;
};

struct HostModel {
    uint32_t vert_base; // offsets into the concatenated vertex tables
    uint32_t face_base;
    uint32_t n_verts, n_faces;
    bool has_normals, has_colours;
};

struct HostLight {
    int geometry; // 0 point, 1 area sphere
    D3 v;
    double radius;
    double colour[3];
    double intensity;
    uint32_t prim; // prim id of the light's sphere, PATHS_MISS for point lights
};

// ---- reference-order binary tree (src/bvh.rs:23-70), children boxes pulled up into the parent
struct RefNode {
    double lbox[6]; // left child: min xyz, max xyz
    double rbox[6];
    int32_t left, right; // >= 0 inner node index; < 0 leaf: ~prim
};
struct RefTree {
    std::vector<RefNode> nodes;
    int32_t root = 0; // >= 0 inner node; < 0 single leaf
    double root_box[6];
    uint32_t depth = 0;
    uint32_t n_nodes_total = 0; // 2N-1
};

// ---- production tree: 8-wide (or 2-wide) nodes over a leaf-ordered primitive stream
struct FastNode2 { // 64 B, both children boxes in the parent
    float c0[6];   // c0 lo.x hi.x lo.y hi.y lo.z hi.z is rearranged at upload; kept simple here
    float c1[6];
    int32_t i0, i1; // >= 0 inner; < 0 leaf code: ~((first << 3) | (count - 1))
    int32_t pad[2];
};
struct FastNode4 { // 128 B = one cache line: child boxes as SoA rows, then the four child codes
    float lox[4], hix[4], loy[4], hiy[4], loz[4], hiz[4];
    int32_t child[4]; // >= 0 inner; < 0 leaf code; empty slots carry an inverted (never hit) box
    int32_t pad[4];
};
struct FastTree {
    std::vector<FastNode2> nodes;
    int32_t root = 0;
    float root_box[6];
    std::vector<uint32_t> leaf_prims; // prim ids in leaf order
    uint32_t depth = 0;
    uint32_t n_giants = 0; // primitives kept out of the SAH build and chained above its root (bvh_fast.cpp)
    // the same tree collapsed to four-wide nodes (every other level of the binary tree is folded into its parent)
    std::vector<FastNode4> nodes4;
    int32_t root4 = 0;
    uint32_t depth4 = 0;
};

struct HostScene {
    std::vector<HostPrim> prims;
    std::vector<HostObject> objects;
    std::vector<HostModel> models;
    std::vector<HostLight> lights;
    // concatenated per-model tables
    std::vector<D3> vertex_normals; // un-normalised means (src/model.rs:194-224); NaN where a vertex has no face
    std::vector<double> vertex_colours; // rgb
    std::vector<uint32_t> faces;       // 3 per face, model-local vertex indices
    int sky_type;
    double sky_c0[3], sky_c1[3];
    uint32_t sky_w = 0, sky_h = 0;
    std::vector<float> sky_texels;
    // importance tables of the Hdri sky (extension ENV_NEE): row marginal CDF [h + 1], per-row conditional CDFs [h][w + 1]
    std::vector<float> env_marginal, env_conditional;
    uint32_t n_tris = 0, n_spheres = 0, hidden = 0;
    D3 bounds_lo, bounds_hi;
    int auto_precision = 32;
};

// desc -> flattened scene.  Returns false and fills err on failure.
bool build_host_scene(const paths_scene_desc &desc, HostScene &out, std::string &err);
// src/bvh.rs:160-206 (AAC), deterministic restatement
bool build_ref_tree(const HostScene &s, RefTree &out, std::string &err);
// binned-SAH production tree
bool build_fast_tree(const HostScene &s, FastTree &out, std::string &err);

uint64_t morton_code(uint16_t num_bits, uint16_t x, uint16_t y, uint16_t z); // src/bvh.rs:373-384
uint32_t ccrf(uint32_t x);                                                    // src/bvh.rs:154-158

} // namespace paths
