// renderer.cuh -- host orchestration of the wavefront pipeline for one GPU; the device-side replacement of
// Renderer (src/renderer.rs:33-141) + Worker (src/worker.rs:43-87) + Estimator (src/pixels.rs:33-80).
// Templated on the arithmetic type; instantiated once per translation unit (kernels_f32.cu / kernels_f64.cu).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "nccl_dyn.h"
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

// ------------------------------------------------------------------ which traversal kernels a scene runs on
// ONE selection shared by RendererImpl (the render) and the per-ray seam (paths_trace_closest / paths_trace_occluded with precision
// 32), so that a fixed-ray test drives exactly the kernels, node layout and grids a render of the same scene launches.
#ifndef PATHS_MIXED_WIDE4
#define PATHS_MIXED_WIDE4 1 /* mixed mode: four-wide nodes for bounce and visibility rays of scenes above 64k primitives */
#endif
struct TraceConfig {
    int sm_count = 148;
    int variant = 3;    // bounce and visibility rays: 0 plain batches (k_extend / k_shadow), 3 two-wide lane refill, 16 four-wide
    int variant_b0 = 1; // camera segment: 1 = k_camera_extend (raygen fused with a plain-batch closest hit), 0/3/16 = k_raygen + that kernel
    int grid_trace = 0, grid_camera = 0, grid4 = 0, grid_camera4 = 0;
    int grid_mixed = 0; // lane-refill kernels of mixed mode (f64 path state over the two-wide tree)
    int grid_mixed4 = 0; // the same over the four-wide nodes (scenes above 64k primitives); 0 = not available
    bool camera_wide4 = false; // camera segment over the four-wide nodes (scenes above 64k primitives)
    size_t smem4 = 0;
};

// stack overflows of the two-wide kernels (dev_trace.cuh) since the library was loaded, on the current device
static inline paths_status read_stack_overflows(unsigned long long *out, unsigned long long *asserts, std::string &err) {
    PATHS_CU(cudaMemcpyFromSymbol(out, g_stack_overflows, sizeof(unsigned long long)));
    PATHS_CU(cudaMemcpyFromSymbol(asserts, g_assert_failures, sizeof(unsigned long long)));
    return PATHS_OK;
}

template <typename R> paths_status configure_trace(const SceneDevice *sd, TraceConfig &tc, std::string &err) {
    cudaDeviceProp prop;
    PATHS_CU(cudaGetDeviceProperties(&prop, sd->device));
    tc.sm_count = prop.multiProcessorCount;
    // two-wide nodes + local stack for small (L1-resident) scenes, four-wide nodes + shared-memory stack above 64k primitives
    // (measured: teapot 7.5 vs 7.8 ms, dragon 25.4 vs 24.4 ms per 16 spp; profiles/r01_*)
    tc.variant = sd->view.n_prims > 65536u ? 16 : 3;
    tc.variant_b0 = 1;
    // measured and NOT adopted (profiles/r02p_cam4_ab.txt): the camera segment over the four-wide nodes is slower than over the two-wide
    // ones even on the large scenes (dragon -1.5 %, environment.yml -10 % whole frame): coherent primary rays keep a warp together on the
    // narrow nodes, and the wide kernel pays its sorting network and shared-memory stack for nothing.  Kept behind a build switch.
#ifndef PATHS_CAMERA_WIDE4_DEFAULT
#define PATHS_CAMERA_WIDE4_DEFAULT 0
#endif
    tc.camera_wide4 = PATHS_CAMERA_WIDE4_DEFAULT && tc.variant == 16;
#ifdef PATHS_DEV_SWITCHES /* development builds only (make EXTRA=-DPATHS_DEV_SWITCHES): a stale variable must not change a release run */
    if (const char *v = getenv("PATHS_EXTEND_VARIANT")) {
        int want = atoi(v);
        if (want == 0 || want == 3 || want == 16) tc.variant = want;
    }
    if (const char *v = getenv("PATHS_EXTEND_B0")) tc.variant_b0 = atoi(v);
    if (const char *v = getenv("PATHS_CAMERA_WIDE4")) tc.camera_wide4 = atoi(v) != 0 && tc.variant == 16;
#endif
    int bps = 0;
    if constexpr (IsF64<R>::value) {
        PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (k_extend<R, false>), 128, 0));
        tc.grid_trace = tc.sm_count * std::max(bps, 1);
        PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (PATHS_K_EXTEND2M), 128, 0));
        tc.grid_mixed = tc.sm_count * std::max(bps, 1);
        // mixed mode over the four-wide nodes: same size rule and shared-memory stack as the f32 kernels
        tc.smem4 = (size_t)(3 * sd->view.fast4_depth + 1) * 128 * sizeof(int32_t);
        if (PATHS_MIXED_WIDE4 && sd->view.n_prims > 65536u && sd->view.fast4_nodes && tc.smem4 <= 200 * 1024) {
            PATHS_CU(cudaFuncSetAttribute(PATHS_K_EXTEND4M, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc.smem4));
            PATHS_CU(cudaFuncSetAttribute(PATHS_K_SHADOW4M, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc.smem4));
            PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, PATHS_K_EXTEND4M, 128, tc.smem4));
            tc.grid_mixed4 = tc.sm_count * std::max(bps, 1);
        } else tc.smem4 = 0;
        tc.variant = 0;
        tc.variant_b0 = 0;
    } else {
        PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (PATHS_K_EXTEND2), 128, 0));
        tc.grid_trace = tc.sm_count * std::max(bps, 1);
        PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (k_camera_extend<float, false>), 128, 0));
        tc.grid_camera = tc.sm_count * std::max(bps, 1);
        // four-wide kernels: stack of 3 * depth + 1 levels per thread in shared memory
        tc.smem4 = (size_t)(3 * sd->view.fast4_depth + 1) * 128 * sizeof(int32_t);
        if (tc.smem4 > 200 * 1024 || !sd->view.fast4_nodes) {
            if (tc.variant == 16) tc.variant = 3; // pathological tree: the two-wide kernels keep their stack in local memory
            tc.smem4 = 0;
            tc.camera_wide4 = false;
        } else {
            PATHS_CU(cudaFuncSetAttribute(PATHS_K_SHADOW4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc.smem4));
            PATHS_CU(cudaFuncSetAttribute(PATHS_K_EXTEND4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc.smem4));
            PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, PATHS_K_EXTEND4, 128, tc.smem4));
            tc.grid4 = tc.sm_count * std::max(bps, 1);
            PATHS_CU(cudaFuncSetAttribute((k_camera_extend<float, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc.smem4));
            PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (k_camera_extend<float, true>), 128, tc.smem4));
            tc.grid_camera4 = tc.sm_count * std::max(bps, 1);
        }
    }
    return PATHS_OK;
}

// ------------------------------------------------------------------ tile-shard combine: owned pixels <-> dense send buffers
template <typename R4> __global__ void k_pack_pixels(const R4 *sums, const uint32_t *pixels, uint32_t n, R4 *out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = sums[pixels[i]];
}
template <typename R4> __global__ void k_scatter_pixels(const R4 *in, const uint32_t *pixels, uint32_t n, R4 *frame) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) frame[pixels[i]] = in[i];
}

// ------------------------------------------------------------------ batch queries
template <typename R, bool MIXED, bool WIDE4 = false>
__global__ void k_trace_closest(DevScene sc, const double *o, const double *d, uint32_t n, uint32_t *prim, double *t, double *bary) {
    extern __shared__ int32_t smem_stack[];
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
        HitF h = WIDE4 ? trace_plain4(sc, (float)ox, (float)oy, (float)oz, (float)dx, (float)dy, (float)dz, smem_stack + threadIdx.x, blockDim.x)
                       : trace_fast<false>(sc, (float)ox, (float)oy, (float)oz, (float)dx, (float)dy, (float)dz, __int_as_float(0x7f800000), 0xFFFFFFFFu);
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

// ---- the per-ray seam over the PRODUCTION kernels: host rays are staged into a PathState exactly like the one a render pass keeps,
// the kernels a render would launch for this scene (TraceConfig) run over it, and the hit / radiance records are read back.
template <typename R4> __global__ void k_stage_rays(const double *o, const double *d, uint32_t n, R4 *ray_pairs) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typedef decltype(ray_pairs->x) S;
    ray_pairs[2 * i] = mk4(S(o[3 * i]), S(o[3 * i + 1]), S(o[3 * i + 2]), bits_to_real(i, S(0))); // {origin | slot}
    ray_pairs[2 * i + 1] = mk4(S(d[3 * i]), S(d[3 * i + 1]), S(d[3 * i + 2]), S(0));
}
template <typename R4> __global__ void k_unstage_hits(const R4 *hit, DevScene sc, uint32_t n, uint32_t *prim, double *t, double *bary) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typedef decltype(hit->x) S;
    R4 h = hit[i];
    uint32_t p = real_to_bits(h.x);
    bool is_hit = p != 0xFFFFFFFFu;
    prim[i] = p;
    t[i] = is_hit ? (double)h.y : (double)__int_as_float(0x7f800000);
    if (bary) {
        bool tri = is_hit && (sc.prim_meta[p].x & PATHS_PM_TRI);
        bary[3 * i] = tri ? (double)h.z : 0.0;
        bary[3 * i + 1] = tri ? (double)h.w : 0.0;
        bary[3 * i + 2] = tri ? (double)(S(1.0) - h.z - h.w) : 0.0;
    }
}
// visibility records as k_shade writes them: origin | light, direction | slot, contribution; slot i's radiance record starts at zero
template <typename R4> __global__ void k_stage_shadow(const double *o, const double *d, const uint32_t *light, uint32_t n, R4 *sh_o, R4 *sh_d, R4 *sh_c,
                                                      R4 *col) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typedef decltype(sh_o->x) S;
    sh_o[i] = mk4(S(o[3 * i]), S(o[3 * i + 1]), S(o[3 * i + 2]), bits_to_real(light ? light[i] : PATHS_LIGHT_ENV, S(0)));
    sh_d[i] = mk4(S(d[3 * i]), S(d[3 * i + 1]), S(d[3 * i + 2]), bits_to_real(i, S(0)));
    sh_c[i] = mk4(S(1), S(2), S(4), S(0));
    col[i] = mk4(S(0), S(0), S(0), S(0));
}
template <typename R4> __global__ void k_unstage_shadow(const R4 *col, uint32_t n, uint8_t *occluded) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typedef decltype(col->x) S;
    R4 c = col[i];
    // an unoccluded record adds its contribution (1, 2, 4) to the slot exactly once; anything else is reported as code 2
    occluded[i] = (c.x == S(0) && c.y == S(0) && c.z == S(0)) ? 1 : ((c.x == S(1) && c.y == S(2) && c.z == S(4)) ? 0 : 2);
}

struct DevScratch { // frees what a batch call allocated, whatever way it returns
    std::vector<void *> ptrs;
    ~DevScratch() {
        for (void *p : ptrs) cudaFree(p);
    }
    template <typename T> cudaError_t get(T **p, size_t count) {
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(q);
        *p = (T *)q;
        return e;
    }
};

// kernel: 0 = the selection a render makes for bounce rays; PATHS_TRACE_CAMERA / _TWO_WIDE / _FOUR_WIDE / _PLAIN force one
template <typename R, bool MIXED = false>
paths_status trace_closest_impl(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                                int kernel, std::string &err) {
    typedef typename Vec4T<R>::type R4;
    if (n == 0) return PATHS_OK;
    if (n > 0x3FFFFFFFu) {
        err = "too many rays in one batch";
        return PATHS_ERR_INVALID_ARG;
    }
    PATHS_CU(cudaSetDevice(sd->device));
    DevScratch mem;
    double *d_o = nullptr, *d_d = nullptr, *d_t = nullptr, *d_b = nullptr;
    uint32_t *d_p = nullptr, *ctl = nullptr;
    PATHS_CU(mem.get(&d_o, n * 3));
    PATHS_CU(mem.get(&d_d, n * 3));
    PATHS_CU(mem.get(&d_t, n));
    PATHS_CU(mem.get(&d_p, n));
    if (bary) PATHS_CU(mem.get(&d_b, n * 3));
    PATHS_CU(cudaMemcpy(d_o, o, n * 24, cudaMemcpyHostToDevice));
    PATHS_CU(cudaMemcpy(d_d, d, n * 24, cudaMemcpyHostToDevice));
    const unsigned blocks = (unsigned)((n + 127) / 128);
    bool staged = false;
    if constexpr (!IsF64<R>::value) {
        if (kernel != PATHS_TRACE_CAMERA) { // the bounce-ray kernels of a render pass, over a staged path state
            TraceConfig tc;
            paths_status st = configure_trace<R>(sd, tc, err);
            if (st != PATHS_OK) return st;
            int variant = kernel == PATHS_TRACE_TWO_WIDE ? 3 : kernel == PATHS_TRACE_FOUR_WIDE ? 16 : kernel == PATHS_TRACE_PLAIN ? 0 : tc.variant;
            if (variant == 16 && tc.grid4 == 0) {
                err = "the four-wide tree of this scene is not staged (or too deep for its shared-memory stack)";
                return PATHS_ERR_UNSUPPORTED;
            }
            PathState<R4> ps;
            std::memset(&ps, 0, sizeof(ps));
            PATHS_CU(mem.get(&ps.ray[0], 2 * n));
            PATHS_CU(mem.get(&ps.hit, n));
            PATHS_CU(mem.get(&ctl, 2));
            const uint32_t host_ctl[2] = {(uint32_t)n, 0u}; // queue length, fetch cursor
            PATHS_CU(cudaMemcpy(ctl, host_ctl, sizeof(host_ctl), cudaMemcpyHostToDevice));
            k_stage_rays<R4><<<blocks, 128>>>(d_o, d_d, (uint32_t)n, ps.ray[0]);
            if (variant == 16) PATHS_K_EXTEND4<<<tc.grid4, 128, tc.smem4>>>(sd->view, ps.ray[0], ps.hit, ctl, ctl + 1);
            else if (variant == 3) PATHS_K_EXTEND2<<<tc.grid_trace, 128>>>(sd->view, ps.ray[0], ps.hit, ctl, ctl + 1);
            else k_extend<R, false><<<tc.grid_trace, 128>>>(sd->view, ps.ray[0], ps.hit, ctl, ctl + 1);
            k_unstage_hits<R4><<<blocks, 128>>>(ps.hit, sd->view, (uint32_t)n, d_p, d_t, d_b);
            staged = true;
        }
    }
    if constexpr (IsF64<R>::value && MIXED) { // mixed mode: the lane-refill kernel a render launches for its bounce rays
        TraceConfig tc;
        paths_status st = configure_trace<R>(sd, tc, err);
        if (st != PATHS_OK) return st;
        PathState<R4> ps;
        std::memset(&ps, 0, sizeof(ps));
        PATHS_CU(mem.get(&ps.ray[0], 2 * n));
        PATHS_CU(mem.get(&ps.hit, n));
        PATHS_CU(mem.get(&ctl, 2));
        const uint32_t host_ctl[2] = {(uint32_t)n, 0u}; // queue length, fetch cursor
        PATHS_CU(cudaMemcpy(ctl, host_ctl, sizeof(host_ctl), cudaMemcpyHostToDevice));
        k_stage_rays<R4><<<blocks, 128>>>(d_o, d_d, (uint32_t)n, ps.ray[0]);
        if (tc.grid_mixed4 > 0) PATHS_K_EXTEND4M<<<tc.grid_mixed4, 128, tc.smem4>>>(sd->view, ps.ray[0], ps.hit, ctl, ctl + 1);
        else PATHS_K_EXTEND2M<<<tc.grid_mixed, 128>>>(sd->view, ps.ray[0], ps.hit, ctl, ctl + 1);
        k_unstage_hits<R4><<<blocks, 128>>>(ps.hit, sd->view, (uint32_t)n, d_p, d_t, d_b);
        staged = true;
    }
    if (!staged) {
        bool wide = false;
        if constexpr (!IsF64<R>::value) { // PATHS_TRACE_CAMERA: the routine the fused camera kernel runs for this scene
            TraceConfig tc;
            paths_status st = configure_trace<R>(sd, tc, err);
            if (st != PATHS_OK) return st;
            if (tc.camera_wide4 && tc.grid_camera4 > 0) {
                PATHS_CU(cudaFuncSetAttribute((k_trace_closest<R, false, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc.smem4));
                k_trace_closest<R, false, true><<<blocks, 128, tc.smem4>>>(sd->view, d_o, d_d, (uint32_t)n, d_p, d_t, d_b);
                wide = true;
            }
        }
        if (!wide) k_trace_closest<R, MIXED><<<blocks, 128>>>(sd->view, d_o, d_d, (uint32_t)n, d_p, d_t, d_b);
    }
    PATHS_CU(cudaGetLastError());
    PATHS_CU(cudaDeviceSynchronize());
    PATHS_CU(cudaMemcpy(prim, d_p, n * 4, cudaMemcpyDeviceToHost));
    PATHS_CU(cudaMemcpy(t, d_t, n * 8, cudaMemcpyDeviceToHost));
    if (bary) PATHS_CU(cudaMemcpy(bary, d_b, n * 24, cudaMemcpyDeviceToHost));
    return PATHS_OK;
}

// The shadow query of trace.rs:61-66 over a batch of host rays, through the visibility kernel a render launches for this scene and
// arithmetic type (k_shadow_refill / k_shadow4 in f32, k_shadow in f64 and mixed mode).  light[i] = index of the light the ray was
// sampled toward (its own primitive does not occlude it), or null: nothing is exempt.  occluded[i]: 1 blocked, 0 clear.
template <typename R, bool MIXED = false>
paths_status trace_occluded_impl(const SceneDevice *sd, const double *o, const double *d, const uint32_t *light, size_t n, uint8_t *occluded,
                                 int kernel, std::string &err) {
    typedef typename Vec4T<R>::type R4;
    if (n == 0) return PATHS_OK;
    if (n > 0x3FFFFFFFu) {
        err = "too many rays in one batch";
        return PATHS_ERR_INVALID_ARG;
    }
    PATHS_CU(cudaSetDevice(sd->device));
    TraceConfig tc;
    paths_status st = configure_trace<R>(sd, tc, err);
    if (st != PATHS_OK) return st;
    int variant = kernel == PATHS_TRACE_TWO_WIDE ? 3 : kernel == PATHS_TRACE_FOUR_WIDE ? 16 : kernel == PATHS_TRACE_PLAIN ? 0 : tc.variant;
    if (IsF64<R>::value) variant = 0;
    if (variant == 16 && tc.grid4 == 0) {
        err = "the four-wide tree of this scene is not staged (or too deep for its shared-memory stack)";
        return PATHS_ERR_UNSUPPORTED;
    }
    DevScratch mem;
    double *d_o = nullptr, *d_d = nullptr;
    uint32_t *d_l = nullptr, *ctl = nullptr;
    uint8_t *d_occ = nullptr;
    unsigned long long *totals = nullptr;
    PathState<R4> ps;
    std::memset(&ps, 0, sizeof(ps));
    PATHS_CU(mem.get(&d_o, n * 3));
    PATHS_CU(mem.get(&d_d, n * 3));
    if (light) PATHS_CU(mem.get(&d_l, n));
    PATHS_CU(mem.get(&d_occ, n));
    PATHS_CU(mem.get(&ps.sh_o, n));
    PATHS_CU(mem.get(&ps.sh_d, n));
    PATHS_CU(mem.get(&ps.sh_c, n));
    PATHS_CU(mem.get(&ps.col, n));
    PATHS_CU(mem.get(&ctl, 2));
    PATHS_CU(mem.get(&totals, 4));
    PATHS_CU(cudaMemset(totals, 0, 4 * sizeof(unsigned long long)));
    const uint32_t host_ctl[2] = {(uint32_t)n, 0u};
    PATHS_CU(cudaMemcpy(ctl, host_ctl, sizeof(host_ctl), cudaMemcpyHostToDevice));
    PATHS_CU(cudaMemcpy(d_o, o, n * 24, cudaMemcpyHostToDevice));
    PATHS_CU(cudaMemcpy(d_d, d, n * 24, cudaMemcpyHostToDevice));
    if (light) PATHS_CU(cudaMemcpy(d_l, light, n * 4, cudaMemcpyHostToDevice));
    const unsigned blocks = (unsigned)((n + 127) / 128);
    k_stage_shadow<R4><<<blocks, 128>>>(d_o, d_d, d_l, (uint32_t)n, ps.sh_o, ps.sh_d, ps.sh_c, ps.col);
    bool launched = false;
    if constexpr (!IsF64<R>::value) {
        if (variant == 16) {
            PATHS_K_SHADOW4<<<tc.grid4, 128, tc.smem4>>>(sd->view, ps, ctl, ctl + 1, totals);
            launched = true;
        } else if (variant == 3) {
            PATHS_K_SHADOW2<<<tc.grid_trace, 128>>>(sd->view, ps, ctl, ctl + 1, totals);
            launched = true;
        }
    }
    if constexpr (IsF64<R>::value && MIXED) {
        if (tc.grid_mixed4 > 0) PATHS_K_SHADOW4M<<<tc.grid_mixed4, 128, tc.smem4>>>(sd->view, ps, ctl, ctl + 1, totals);
        else PATHS_K_SHADOW2M<<<tc.grid_mixed, 128>>>(sd->view, ps, ctl, ctl + 1, totals);
        launched = true;
    }
    if (!launched) k_shadow<R, MIXED><<<tc.grid_trace, 128>>>(sd->view, ps, ctl, ctl + 1, totals);
    k_unstage_shadow<R4><<<blocks, 128>>>(ps.col, (uint32_t)n, d_occ);
    PATHS_CU(cudaGetLastError());
    PATHS_CU(cudaDeviceSynchronize());
    PATHS_CU(cudaMemcpy(occluded, d_occ, n, cudaMemcpyDeviceToHost));
    return PATHS_OK;
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
