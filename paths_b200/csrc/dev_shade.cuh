// dev_shade.cuh -- camera, materials, lights and sky of the reference, as device functions templated on the arithmetic type.
#pragma once
#include "dev_math.cuh"
#include "dev_rng.cuh"
#include "dev_scene.h"

namespace paths {

// ------------------------------------------------------------------ camera (src/camera.rs:41-94)
template <typename R> __device__ __forceinline__ void camera_ray(const DevCamera &cam, uint32_t px, uint32_t py, R sqx, R sqy, R dkx, R dky,
                                                                  V3<R> &origin, V3<R> &direction, R &weight) {
    uint32_t x = cam.width - px - 1; // :56-57
    uint32_t y = cam.height - py - 1;
    R f = R(cam.focal_length), v = R(cam.distance_from_lens);
    R p = (f * v) / (v - f); // :65-67
    R x_scale = R(cam.sensor_width) / R(cam.width);
    R y_scale = R(cam.sensor_height) / R(cam.height);
    R image_x = R(x) - R(cam.width) / R(2.0) + sqx; // :73-76
    R image_y = R(cam.height) / R(2.0) - R(y) - sqy;
    V3<R> k = mk3<R>(image_x * x_scale, image_y * y_scale, -v);
    R aperture_radius = f / R(cam.aperture); // :41-45
    V3<R> l = mk3<R>(dkx * aperture_radius, dky * aperture_radius, R(0));
    V3<R> dir = neg((k * (p / v)) + l); // :83
    V3<R> norm_dir = normed(dir);
    V3<R> loc = mk3<R>(R(cam.location[0]), R(cam.location[1]), R(cam.location[2]));
    origin = mat_apply<R>(cam.rot, l) + loc; // :87-88
    direction = mat_apply<R>(cam.rot, norm_dir);
    weight = norm_dir.z; // :91
}

// worker.rs:58-87 with the per-request CMJ seeds drawn from the counter-based generator
template <typename R> __device__ __forceinline__ void pixel_sample_ray(const DevCamera &cam, uint32_t k0, uint32_t k1, uint32_t pixel,
                                                                        uint32_t S, V3<R> &o, V3<R> &d, R &weight) {
    uint32_t m, n, s, pattern;
    sample_schedule(S, m, n, s, pattern);
    U4 seeds = philox4x32_10(pixel, pattern, 0u, PATHS_TAG_CMJ, k0, k1);
    R sqx, sqy, dkx, dky;
    cmj<R>(s, m, n, seeds.v[0], sqx, sqy);
    cmj_disk<R>(s, m, n, seeds.v[1], dkx, dky);
    camera_ray<R>(cam, pixel % cam.width, pixel / cam.width, sqx, sqy, dkx, dky, o, d, weight);
}

// ------------------------------------------------------------------ materials (src/material.rs)
template <typename R> struct MatR {
    int type;
    V3<R> albedo;
    R r0, metalness, roughness;
    bool ggx; // extension PATHS_EXT_GGX: CookTorrance evaluated with the GGX distribution instead of Beckmann
};

template <typename R> __device__ __forceinline__ V3<R> reflect(V3<R> vector, V3<R> normal) { // :249-251
    return normed((normal * dot(normal, vector) * R(2.0)) - vector);
}
// geom.rs:10-24 (note y = 1 - u, not sqrt(1 - u): kept)
template <typename R> __device__ __forceinline__ V3<R> lambert_direction(V3<R> normal, R u, R v) { // material.rs:224-231
    R r = sqrt_(u);
    R sn, cs;
    sincos_2pi(v, &sn, &cs); // theta = 2 pi v
    V3<R> rd = mk3<R>(r * cs, R(1.0) - u, r * sn);
    return normed(to_world(rd, normal));
}
template <typename R> __device__ __forceinline__ R lambert_pdf(V3<R> vec_in, V3<R> normal) { return dot(normal, neg(vec_in)) / Consts<R>::pi(); } // :220-222
template <typename R> __device__ __forceinline__ V3<R> lambert_brdf(V3<R> albedo, V3<R> vec_in, V3<R> normal) {                                   // :237-239
    return albedo * dot(normal, neg(vec_in)) / Consts<R>::pi();
}
template <typename R> __device__ __forceinline__ R schlick(R r0, V3<R> vec_out, V3<R> normal) { // :303-305
    R cos_theta = dot(vec_out, normal);
    return r0 + (R(1.0) - r0) * pow5_(R(1.0) - cos_theta);
}
template <typename R> __device__ __forceinline__ R beckmann_ndf(R m, V3<R> n, V3<R> h) { // :436-447
    R alpha = acos_(dot(h, n));
    R cos_alpha = cos(alpha);
    R tan_alpha = tan_(alpha);
    R e = exp_(R(-1.0) * (tan_alpha * tan_alpha) / (m * m));
    R d0 = e / (Consts<R>::pi() * m * m * pow_(cos_alpha, R(4.0)));
    return fmax_(R(0.0), d0 * dot(n, h));
}
// Production (f32): h and n are normalised with rsqrtf (about 2 ulp each), so their dot product can exceed 1 at the centre of a
// highlight and acosf would return NaN.  Same function without the inverse trigonometry: cos(alpha) = c clamped to [-1, 1],
// tan^2(alpha) = (1 - c^2) / c^2.
template <> __device__ __forceinline__ float beckmann_ndf<float>(float m, V3<float> n, V3<float> h) {
    float c = fminf(1.0f, fmaxf(-1.0f, dot(h, n)));
    float c2 = c * c;
    if (!(c2 > 0.0f)) return 0.0f;
    float tan2 = (1.0f - c2) / c2;
    float e = expf(-tan2 / (m * m));
    float d0 = e / (Consts<float>::pi() * m * m * c2 * c2);
    return fmaxf(0.0f, d0 * c);
}
// extension: the GGX / Trowbridge-Reitz distribution in the same D * (n.h) form as ct_ndf above
template <typename R> __device__ __forceinline__ R ggx_ndf(R m, V3<R> n, V3<R> h) {
    R c = dot(n, h);
    R a2 = m * m;
    R q = c * c * (a2 - R(1.0)) + R(1.0);
    return fmax_(R(0.0), a2 / (Consts<R>::pi() * q * q) * c);
}
template <typename R> __device__ __forceinline__ V3<R> basic_brdf(const MatR<R> &b, V3<R> vec_out, V3<R> vec_in, V3<R> normal) {
    const V3<R> black = mk3<R>(R(0), R(0), R(0));
    if (b.type == 0) return lambert_brdf(b.albedo, vec_in, normal);
    if (b.type == 2) { // :360-370
        R r = schlick(b.r0, vec_out, normal);
        V3<R> diffuse = lambert_brdf(b.albedo, vec_in, normal) * (R(1.0) - b.metalness);
        return diffuse * (R(1.0) - r) + black * r;
    }
    if (b.type == 3) { // :503-523
        V3<R> h = normed(vec_out - vec_in);
        R d = b.ggx ? ggx_ndf(b.roughness, normal, h) : beckmann_ndf(b.roughness, normal, h);
        R ndl = dot(normal, neg(vec_in));
        R vdh = dot(vec_out, h);
        R ndh = dot(normal, h);
        R ndv = dot(normal, vec_out);
        R g = fmax_(R(0.0), fmin_(R(1.0), fmin_((R(2.0) * ndh * ndv) / vdh, (R(2.0) * ndh * ndl) / vdh)));
        return b.albedo * (d * g) / (R(4.0) * ndv * ndl);
    }
    return black; // Mirror :268-271
}

template <typename R> struct Material {
    MatR<R> m;
    bool fresnel;
    R fresnel_r0;
    MatR<R> diffuse, specular;
};

template <typename R, typename S> __device__ __forceinline__ Material<R> load_material(const DevObjectT<S> &o) {
    Material<R> mt;
    mt.m.type = o.mat_type;
    mt.m.albedo = mk3<R>(R(o.albedo[0]), R(o.albedo[1]), R(o.albedo[2]));
    mt.m.r0 = R(o.r0);
    mt.m.metalness = R(o.metalness);
    mt.m.roughness = R(o.roughness);
    mt.m.ggx = mt.diffuse.ggx = mt.specular.ggx = false;
    mt.fresnel = o.mat_type == 4;
    if (mt.fresnel) {
        mt.fresnel_r0 = R(o.fresnel_r0);
        mt.diffuse.type = o.d_type;
        mt.diffuse.albedo = mk3<R>(R(o.d_albedo[0]), R(o.d_albedo[1]), R(o.d_albedo[2]));
        mt.diffuse.r0 = R(o.d_r0); mt.diffuse.metalness = R(o.d_metal); mt.diffuse.roughness = R(o.d_rough);
        mt.specular.type = o.s_type;
        mt.specular.albedo = mk3<R>(R(o.s_albedo[0]), R(o.s_albedo[1]), R(o.s_albedo[2]));
        mt.specular.r0 = R(o.s_r0); mt.specular.metalness = R(o.s_metal); mt.specular.roughness = R(o.s_rough);
    }
    return mt;
}

template <typename R> __device__ __forceinline__ V3<R> material_brdf(const Material<R> &m, V3<R> vec_out, V3<R> vec_in, V3<R> normal) {
    if (m.fresnel) { // :414-421
        R r = schlick(m.fresnel_r0, vec_out, normal);
        return basic_brdf(m.diffuse, vec_out, vec_in, normal) * (R(1.0) - r) + basic_brdf(m.specular, vec_out, vec_in, normal) * r;
    }
    return basic_brdf(m.m, vec_out, vec_in, normal);
}

// material.rs:81-88: Lambertian / Mirror / Gloss only.  us = {gloss choice, hemisphere u, hemisphere v}.
template <typename R> __device__ __forceinline__ bool material_sample(const Material<R> &mt, V3<R> vec_out, V3<R> normal, R u0, R u1, R u2,
                                                                       V3<R> &direction, R &pdf, V3<R> &brdf, bool &is_specular) {
    if (mt.fresnel) return false;
    const MatR<R> &b = mt.m;
    const V3<R> white = mk3<R>(R(1), R(1), R(1));
    if (b.type == 0) { // :211-216
        direction = lambert_direction(normal, u1, u2);
        pdf = lambert_pdf(neg(direction), normal);
        brdf = lambert_brdf(b.albedo, neg(direction), normal);
        is_specular = false;
        return true;
    }
    if (b.type == 1) { // :253-255
        direction = reflect(vec_out, normal);
        pdf = R(1.0);
        brdf = white;
        is_specular = true;
        return true;
    }
    if (b.type == 2) { // :302-325
        R r = schlick(b.r0, vec_out, normal);
        R chance = b.r0 > R(0.5) ? r : R(0.5);
        is_specular = u0 <= chance;
        if (is_specular) {
            direction = reflect(vec_out, normal);
            V3<R> br = b.albedo * b.metalness + white * (R(1.0) - b.metalness);
            pdf = R(1.0) * chance;
            brdf = br * r;
        } else {
            direction = lambert_direction(normal, u1, u2);
            V3<R> vec_in = neg(direction);
            R p = lambert_pdf(vec_in, normal);
            V3<R> br = lambert_brdf(b.albedo, vec_in, normal) * (R(1.0) - b.metalness);
            pdf = p * (R(1.0) - chance);
            brdf = br * (R(1.0) - r);
        }
        return true;
    }
    return false;
}

// ------------------------------------------------------------------ extensions without reference behaviour (PATHS_EXT_*)
// Microfacet sampling.  The reference evaluates CookTorrance / Fresnel combinations (brdf(), material.rs:414-421,503-523) but its
// Material::sample panics for them (material.rs:81-88).  Here: sample a Beckmann (or GGX) half vector h with density D(h)(n.h),
// reflect vec_out about it, pdf(dir) = D(h)(n.h) / (4 vec_out.h); the weight is the REFERENCE's brdf() over that pdf, so the
// estimator integrates exactly the function the reference's NEE term uses.
template <typename R> __device__ __forceinline__ bool cook_torrance_sample(const MatR<R> &b, V3<R> vec_out, V3<R> normal, R u1, R u2, bool cosine,
                                                                            V3<R> &direction, R &pdf, V3<R> &brdf) {
    R sn, cs;
    sincos_2pi(u2, &sn, &cs);
    if (cosine) { // plain cosine lobe (cross-check estimator)
        R r = sqrt_(u1), y = sqrt_(fmax_(R(0.0), R(1.0) - u1));
        direction = normed(to_world(mk3<R>(r * cs, y, r * sn), normal));
        pdf = dot(normal, direction) / Consts<R>::pi();
    } else {
        R m = b.roughness;
        R tan2 = b.ggx ? m * m * u1 / fmax_(R(1e-12), R(1.0) - u1) : -m * m * log(fmax_(R(1e-30), R(1.0) - u1));
        R cos_t = R(1.0) / sqrt_(R(1.0) + tan2);
        R sin_t = sqrt_(fmax_(R(0.0), R(1.0) - cos_t * cos_t));
        V3<R> h = normed(to_world(mk3<R>(sin_t * cs, cos_t, sin_t * sn), normal));
        R vdh = dot(vec_out, h);
        if (!(vdh > R(0))) return false;
        direction = normed(h * (R(2.0) * vdh) - vec_out);
        R d = b.ggx ? ggx_ndf(m, normal, h) : beckmann_ndf(m, normal, h);
        pdf = d / (R(4.0) * vdh);
    }
    if (!(dot(normal, direction) > R(0)) || !(pdf > R(0))) return false;
    brdf = basic_brdf(b, vec_out, neg(direction), normal);
    return true;
}

template <typename R> __device__ __forceinline__ bool basic_sample_ext(const MatR<R> &b, V3<R> vec_out, V3<R> normal, R u0, R u1, R u2, bool cosine,
                                                                        V3<R> &direction, R &pdf, V3<R> &brdf, bool &is_specular) {
    if (b.type == 3) {
        is_specular = false;
        return cook_torrance_sample(b, vec_out, normal, u1, u2, cosine, direction, pdf, brdf);
    }
    Material<R> tmp;
    tmp.m = b;
    tmp.fresnel = false;
    return material_sample<R>(tmp, vec_out, normal, u0, u1, u2, direction, pdf, brdf, is_specular);
}

// us = {branch choice, gloss choice, hemisphere / half-vector u, v}
template <typename R> __device__ __forceinline__ bool material_sample_ext(const Material<R> &mt, V3<R> vec_out, V3<R> normal, R uc, R u0, R u1, R u2,
                                                                           bool cosine, V3<R> &direction, R &pdf, V3<R> &brdf, bool &is_specular) {
    if (!mt.fresnel) return basic_sample_ext(mt.m, vec_out, normal, u0, u1, u2, cosine, direction, pdf, brdf, is_specular);
    // one-sample estimator of r f_s + (1 - r) f_d (material.rs:414-421): pick a lobe with probability ps, divide by it
    R r = schlick(mt.fresnel_r0, vec_out, normal);
    R ps = fmin_(R(0.9), fmax_(R(0.1), r));
    if (uc < ps) {
        if (!basic_sample_ext(mt.specular, vec_out, normal, u0, u1, u2, cosine, direction, pdf, brdf, is_specular)) return false;
        brdf = brdf * r;
        pdf = pdf * ps;
    } else {
        if (!basic_sample_ext(mt.diffuse, vec_out, normal, u0, u1, u2, cosine, direction, pdf, brdf, is_specular)) return false;
        brdf = brdf * (R(1.0) - r);
        pdf = pdf * (R(1.0) - ps);
    }
    return true;
}

// Environment importance sampling: picks a texel of the Hdri sky with probability proportional to luminance x solid angle
// (tables built by host_scene.cpp), a uniform point inside it, and returns the LOOKUP direction (what Skybox::ambient_light is
// called with, scene.rs:95-111), its solid-angle density and the texel radiance.
template <typename R> __device__ __forceinline__ void env_sample(const DevScene &sc, R u1, R u2, V3<R> &lookup_dir, R &pdf, V3<R> &radiance) {
    const uint32_t w = sc.sky_w, h = sc.sky_h;
    uint32_t lo = 0, hi = h;
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (sc.env_marginal[mid + 1] <= (float)u1) lo = mid + 1;
        else hi = mid;
    }
    uint32_t y = min(lo, h - 1);
    float m0 = sc.env_marginal[y], m1 = sc.env_marginal[y + 1];
    R fy = fmin_(R(0.999999), fmax_(R(0.0), (u1 - R(m0)) / fmax_(R(1e-20), R(m1 - m0))));
    const float *cond = sc.env_conditional + (size_t)y * (w + 1);
    lo = 0;
    hi = w;
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (cond[mid + 1] <= (float)u2) lo = mid + 1;
        else hi = mid;
    }
    uint32_t x = min(lo, w - 1);
    float c0 = cond[x], c1 = cond[x + 1];
    R fx = fmin_(R(0.999999), fmax_(R(0.0), (u2 - R(c0)) / fmax_(R(1e-20), R(c1 - c0))));
    R lat = Consts<R>::pi() * (R(1.0) - (R(y) + fy) / R(h));
    R lon = Consts<R>::pi() * (R(2.0) * (R(x) + fx) / R(w) - R(1.0));
    R sl, cl, so, co;
    sincos_(lat, &sl, &cl);
    sincos_(lon, &so, &co);
    lookup_dir = mk3<R>(sl * co, cl, sl * so);
    R prob = R(m1 - m0) * R(c1 - c0);
    pdf = prob * R(w) * R(h) / (R(2.0) * Consts<R>::pi() * Consts<R>::pi() * fmax_(R(1e-12), sl));
    size_t pix = (size_t)y * w + x;
    radiance = mk3<R>(R(sc.sky_texels[3 * pix]), R(sc.sky_texels[3 * pix + 1]), R(sc.sky_texels[3 * pix + 2]));
}

// ------------------------------------------------------------------ lights (src/scene.rs:54-59, src/geom.rs:146-173)
template <typename R> __device__ __forceinline__ void light_sample(const DevLight &l, V3<R> from, R u, R v, V3<R> &dir, R &inv_pdf) {
    V3<R> lv = mk3<R>(R(l.v[0]), R(l.v[1]), R(l.v[2]));
    if (l.geometry == 0) {
        dir = lv;
        inv_pdf = R(1.0);
        return;
    }
    R radius = R(l.radius);
    R phi = acos_(R(2.0) * v - R(1.0));
    R sp, cp, st, ct;
    sincos_(phi, &sp, &cp);
    sincos_2pi(u, &st, &ct); // theta = 2 pi u
    V3<R> n = mk3<R>(sp * ct, sp * st, cp);
    V3<R> point = lv + n * radius;
    V3<R> out_vec = from - point;
    dir = normed(out_vec);
    R distance_sq = dot(out_vec, out_vec);
    R area = R(4.0) * Consts<R>::pi() * radius * radius;
    R ip = area * dot(n, dir) / distance_sq;
    inv_pdf = fmax_(R(0.0), ip);
}

// ------------------------------------------------------------------ sky (src/scene.rs:88-113)
__device__ __forceinline__ uint32_t sat_u32(double v) { return (v > 0.0) ? (v >= 4294967295.0 ? 4294967295u : (uint32_t)v) : 0u; }
__device__ __forceinline__ uint32_t sat_u32(float v) { return (v > 0.0f) ? (v >= 4294967040.0f ? 4294967295u : (uint32_t)v) : 0u; }

template <typename R> __device__ __forceinline__ V3<R> sky_light(const DevScene &sc, V3<R> direction) {
    V3<R> c0 = mk3<R>(R(sc.sky_c0[0]), R(sc.sky_c0[1]), R(sc.sky_c0[2]));
    if (sc.sky_type == 0) return c0;
    if (sc.sky_type == 1) {
        V3<R> c1 = mk3<R>(R(sc.sky_c1[0]), R(sc.sky_c1[1]), R(sc.sky_c1[2]));
        R cos_theta = dot(direction, mk3<R>(R(0), R(1), R(0)));
        return c0 * cos_theta + c1 * (R(1.0) - cos_theta);
    }
    R lat = acos_(direction.y);
    R lon = atan2_(direction.z, direction.x);
    R w = R(sc.sky_w), h = R(sc.sky_h);
    R x = w / R(2.0) * (lon / Consts<R>::pi()) + w / R(2.0);
    R y = h * (R(1.0) - lat / Consts<R>::pi());
    uint32_t xp = min(sc.sky_w - 1, sat_u32(x));
    uint32_t yp = min(sc.sky_h - 1, sat_u32(y));
    size_t pix = (size_t)yp * sc.sky_w + xp;
    return mk3<R>(R(sc.sky_texels[3 * pix]), R(sc.sky_texels[3 * pix + 1]), R(sc.sky_texels[3 * pix + 2]));
}

} // namespace paths
