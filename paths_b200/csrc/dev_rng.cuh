// dev_rng.cuh -- per-pixel counter-based random numbers and the reference's correlated multi-jittered sampler.
//
// The reference draws from an unseedable thread_rng (src/trace.rs:106, src/geom.rs:11,149, src/material.rs:310,
// src/scene.rs:201, src/sampling.rs:183).  The B200 path replaces it by Philox4x32-10 keyed by the render seed with the
// counter (pixel, sample, 2*bounce + block, tag): no state to carry between wavefront kernels, and any (pixel, sample) can
// be regenerated on any GPU, which is what makes tile and sample sharding bit-reproducible.
// The CMJ functions restate src/sampling.rs:187-235 (including its `while i > l` permute).
#pragma once
#include "dev_math.cuh"

namespace paths {

#define PATHS_TAG_PATH 0x50415448u /* 'PATH' */
#define PATHS_TAG_CMJ 0x434D4A20u  /* 'CMJ ' */
#define PATHS_TAG_ENV 0x454E5620u  /* 'ENV ' : environment NEE samples (extension) */

struct U4 {
    uint32_t v[4];
};

__host__ __device__ __forceinline__ U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        c1 = (uint32_t)p1;
        c3 = (uint32_t)p0;
        c0 = n0;
        c2 = n2;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    U4 o;
    o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
    return o;
}

// 24-bit uniform in [0,1): exactly representable in f32, so the f32 and f64 instantiations see the same numbers
template <typename R> __device__ __forceinline__ R u01(uint32_t x) { return R(x >> 8) * R(1.0 / 16777216.0); }

// sampling.rs:187-210
__host__ __device__ __forceinline__ uint32_t cmj_permute(uint32_t i, uint32_t l, uint32_t p) {
    uint32_t w = l - 1;
    w |= w >> 1; w |= w >> 2; w |= w >> 4; w |= w >> 8; w |= w >> 16;
    while (i > l) { // sic: never entered for i < l, so the result is the cyclic shift below
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
    return (i + p) % l;
}
// sampling.rs:212-221
__host__ __device__ __forceinline__ uint32_t cmj_hash(uint32_t i, uint32_t p) {
    i ^= p;
    i ^= i >> 17;
    i ^= i >> 10; i *= 0xb36534e5u;
    i ^= i >> 12;
    i ^= i >> 21; i *= 0x93fc4795u;
    i ^= 0xdf6e307fu;
    i ^= i >> 17; i *= (1u | p >> 18);
    return i;
}
template <typename R> __device__ __forceinline__ R cmj_rand_float(uint32_t i, uint32_t p) { return R(cmj_hash(i, p)) * R(1.0 / 4294967808.0); }

// sampling.rs:226-235
template <typename R> __device__ __forceinline__ void cmj_mn(uint32_t s, uint32_t m, uint32_t n, uint32_t p, R &x, R &y) {
    uint32_t ps = cmj_permute(s, m * n, p * 0xa73bd290u);
    R sx = R(cmj_permute(ps % m, m, p * 0xa511e9b3u));
    R sy = R(cmj_permute(ps / m, n, p * 0x63d83595u));
    R jx = cmj_rand_float<R>(s, p * 0xa399d265u);
    R jy = cmj_rand_float<R>(s, p * 0x711ad6a5u);
    x = (R(s % m) + (sy + jx) / R(n)) / R(m);
    y = (R(s / m) + (sx + jy) / R(m)) / R(n);
}
// every sample but a pixel's first comes from a (5, 5) pattern (sample_schedule below): with m and n known at compile time the
// integer remainders and quotients become multiplications (same values; the float divisions stay divisions)
template <typename R> __device__ __forceinline__ void cmj(uint32_t s, uint32_t m, uint32_t n, uint32_t p, R &x, R &y) {
    if (m == 5u && n == 5u) cmj_mn<R>(s, 5u, 5u, p, x, y);
    else cmj_mn<R>(s, m, n, p, x, y);
}
// sampling.rs:250-264
template <typename R> __device__ __forceinline__ void cmj_disk(uint32_t s, uint32_t m, uint32_t n, uint32_t p, R &dx, R &dy) {
    R x, y;
    cmj<R>(s, m, n, p, x, y);
    R r = sqrt_(y);
    R sn, cs;
    sincos_2pi(x, &sn, &cs); // theta = 2 pi x
    dx = r * cs;
    dy = r * sn;
}

// renderer.rs:166-192 + worker.rs:58-87: sample 0 of a pixel comes from a (1,1) pattern, later ones from successive (5,5) patterns
__host__ __device__ __forceinline__ void sample_schedule(uint32_t S, uint32_t &m, uint32_t &n, uint32_t &s, uint32_t &pattern) {
    if (S == 0) { m = 1; n = 1; s = 0; pattern = 0; }
    else { m = 5; n = 5; s = (S - 1) % 25; pattern = 1 + (S - 1) / 25; }
}

} // namespace paths
