// renderer.cuh -- host orchestration of the wavefront pipeline for one GPU; the device-side replacement of
// Renderer (src/renderer.rs:33-141) + Worker (src/worker.rs:43-87) + Estimator (src/pixels.rs:33-80).
// Templated on the arithmetic type; instantiated once per translation unit (kernels_f32.cu / kernels_f64.cu).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "pipeline.cuh"
#include "scene_device.h"

namespace paths {

#define PATHS_CU(call)                                                                        \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            err = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
            return PATHS_ERR_CUDA;                                                            \
        }                                                                                     \
    } while (0)

// ------------------------------------------------------------------ small output kernels
#define PATHS_PREVIEW_GRID 6u /* renderer.rs:13 */
// Estimator::render (pixels.rs:53-79): mean = sum / count; a pixel without samples shows the preview-grid pixel above-left of it
// (x - x % 6, y - y % 6); MeanVec::get of a pixel that has no samples either is 0/0 = NaN, which Colour::to_bytes turns into 0.
template <typename R> __device__ __forceinline__ typename Vec4T<R>::type estimator_pixel(const typename Vec4T<R>::type *sums, uint32_t i, uint32_t W) {
    typename Vec4T<R>::type s = sums[i];
    if (s.w > R(0)) return s;
    uint32_t x = i % W, y = i / W;
    return sums[(x - x % PATHS_PREVIEW_GRID) + (y - y % PATHS_PREVIEW_GRID) * W];
}
template <typename R> __global__ void k_image_f32(const typename Vec4T<R>::type *sums, uint32_t n, uint32_t W, float *rgb) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename Vec4T<R>::type s = estimator_pixel<R>(sums, i, W);
    bool has = s.w > R(0); // the float image reports pixels that nothing covers as 0 rather than NaN
    rgb[3 * i] = has ? (float)(s.x / s.w) : 0.0f;
    rgb[3 * i + 1] = has ? (float)(s.y / s.w) : 0.0f;
    rgb[3 * i + 2] = has ? (float)(s.z / s.w) : 0.0f;
}
// Colour::component_to_byte (colour.rs:62-74); Rust `as u8` saturates and maps NaN to 0
template <typename R> __device__ __forceinline__ uint8_t to_byte(R x) {
    x = pow_(x, R(0.45));
    if (x >= R(1.0)) return 255;
    if (x <= R(0.0)) return 0;
    R v = x * R(256.0);
    if (isnan_(v)) return 0;
    if (v >= R(255.0)) return 255;
    return (uint8_t)v;
}
template <typename R> __global__ void k_image_rgb24(const typename Vec4T<R>::type *sums, uint32_t n, uint32_t W, uint8_t *rgb) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename Vec4T<R>::type s = estimator_pixel<R>(sums, i, W);
    bool has = s.w > R(0);
    rgb[3 * i] = has ? to_byte<R>(s.x / s.w) : 0;
    rgb[3 * i + 1] = has ? to_byte<R>(s.y / s.w) : 0;
    rgb[3 * i + 2] = has ? to_byte<R>(s.z / s.w) : 0;
}
template <typename R> __global__ void k_sums_to_f64(const typename Vec4T<R>::type *sums, uint32_t n, double4 *out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename Vec4T<R>::type s = sums[i];
    out[i] = make_double4((double)s.x, (double)s.y, (double)s.z, (double)s.w);
}
template <typename R> __global__ void k_add_sums(typename Vec4T<R>::type *sums, uint32_t n, const double4 *in) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename Vec4T<R>::type s = sums[i];
    double4 a = in[i];
    s.x += R(a.x); s.y += R(a.y); s.z += R(a.z); s.w += R(a.w);
    sums[i] = s;
}

// ------------------------------------------------------------------ batch queries
template <typename R, bool MIXED>
__global__ void k_trace_closest(DevScene sc, const double *o, const double *d, uint32_t n, uint32_t *prim, double *t, double *bary) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double ox = o[3 * i], oy = o[3 * i + 1], oz = o[3 * i + 2], dx = d[3 * i], dy = d[3 * i + 1], dz = d[3 * i + 2];
    uint32_t p;
    double tt, bx, by;
    if constexpr (IsF64<R>::value) {
        Hit64 h = MIXED ? trace_mixed<false>(sc, mk3(ox, oy, oz), mk3(dx, dy, dz), (double)__int_as_float(0x7f800000), 0xFFFFFFFFu)
                        : trace_ref(sc, mk3(ox, oy, oz), mk3(dx, dy, dz));
        p = h.prim; tt = h.t; bx = h.bx; by = h.by;
    } else {
        HitF h = trace_fast<false>(sc, (float)ox, (float)oy, (float)oz, (float)dx, (float)dy, (float)dz, __int_as_float(0x7f800000), 0xFFFFFFFFu);
        p = h.prim; tt = h.t; bx = h.bx; by = h.by;
    }
    prim[i] = p;
    bool hit = p != 0xFFFFFFFFu;
    t[i] = hit ? tt : (double)__int_as_float(0x7f800000);
    if (bary) {
        bool tri = hit && (sc.prim_meta[p].x & PATHS_PM_TRI);
        bary[3 * i] = tri ? bx : 0.0;
        bary[3 * i + 1] = tri ? by : 0.0;
        bary[3 * i + 2] = tri ? (double)(R(1.0) - R(bx) - R(by)) : 0.0;
    }
}

template <typename R>
__global__ void k_camera_rays(DevCamera cam, uint32_t k0, uint32_t k1, uint32_t sample, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h, double *o,
                              double *d, double *weight) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= w * h) return;
    uint32_t x = x0 + i % w, y = y0 + i / w; // PixelGridIter (worker.rs:170-186)
    V3<R> ro, rd;
    R wt;
    pixel_sample_ray<R>(cam, k0, k1, x + y * cam.width, sample, ro, rd, wt);
    o[3 * i] = ro.x; o[3 * i + 1] = ro.y; o[3 * i + 2] = ro.z;
    d[3 * i] = rd.x; d[3 * i + 1] = rd.y; d[3 * i + 2] = rd.z;
    weight[i] = wt;
}

template <typename R, bool MIXED = false>
paths_status trace_closest_impl(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                                std::string &err) {
    if (n == 0) return PATHS_OK;
    if (n > 0x7FFFFFFFu) {
        err = "too many rays in one batch";
        return PATHS_ERR_INVALID_ARG;
    }
    PATHS_CU(cudaSetDevice(sd->device));
    double *d_o = nullptr, *d_d = nullptr, *d_t = nullptr, *d_b = nullptr;
    uint32_t *d_p = nullptr;
    paths_status st = PATHS_OK;
    auto body = [&]() -> paths_status {
        PATHS_CU(cudaMalloc(&d_o, n * 24));
        PATHS_CU(cudaMalloc(&d_d, n * 24));
        PATHS_CU(cudaMalloc(&d_t, n * 8));
        PATHS_CU(cudaMalloc(&d_p, n * 4));
        if (bary) PATHS_CU(cudaMalloc(&d_b, n * 24));
        PATHS_CU(cudaMemcpy(d_o, o, n * 24, cudaMemcpyHostToDevice));
        PATHS_CU(cudaMemcpy(d_d, d, n * 24, cudaMemcpyHostToDevice));
        k_trace_closest<R, MIXED><<<(unsigned)((n + 127) / 128), 128>>>(sd->view, d_o, d_d, (uint32_t)n, d_p, d_t, d_b);
        PATHS_CU(cudaGetLastError());
        PATHS_CU(cudaDeviceSynchronize());
        PATHS_CU(cudaMemcpy(prim, d_p, n * 4, cudaMemcpyDeviceToHost));
        PATHS_CU(cudaMemcpy(t, d_t, n * 8, cudaMemcpyDeviceToHost));
        if (bary) PATHS_CU(cudaMemcpy(bary, d_b, n * 24, cudaMemcpyDeviceToHost));
        return PATHS_OK;
    };
    st = body();
    cudaFree(d_o); cudaFree(d_d); cudaFree(d_t); cudaFree(d_p); cudaFree(d_b);
    return st;
}

template <typename R>
paths_status camera_rays_impl(const paths_camera &cam, int device, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h,
                              double *o, double *d, double *weight, std::string &err) {
    size_t n = (size_t)w * h;
    if (n == 0) return PATHS_OK;
    PATHS_CU(cudaSetDevice(device));
    double *d_o = nullptr, *d_d = nullptr, *d_w = nullptr;
    auto body = [&]() -> paths_status {
        PATHS_CU(cudaMalloc(&d_o, n * 24));
        PATHS_CU(cudaMalloc(&d_d, n * 24));
        PATHS_CU(cudaMalloc(&d_w, n * 8));
        k_camera_rays<R><<<(unsigned)((n + 127) / 128), 128>>>(make_dev_camera(cam), (uint32_t)seed, (uint32_t)(seed >> 32), sample, x0, y0, w, h, d_o,
                                                               d_d, d_w);
        PATHS_CU(cudaGetLastError());
        PATHS_CU(cudaDeviceSynchronize());
        PATHS_CU(cudaMemcpy(o, d_o, n * 24, cudaMemcpyDeviceToHost));
        PATHS_CU(cudaMemcpy(d, d_d, n * 24, cudaMemcpyDeviceToHost));
        PATHS_CU(cudaMemcpy(weight, d_w, n * 8, cudaMemcpyDeviceToHost));
        return PATHS_OK;
    };
    paths_status st = body();
    cudaFree(d_o); cudaFree(d_d); cudaFree(d_w);
    return st;
}

} // namespace paths
