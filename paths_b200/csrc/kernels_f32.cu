// kernels_f32.cu -- production instantiation of the wavefront pipeline (f32 traversal and shading, f64 spheres).
#include <cstdint>
#include <mutex>

#include "renderer_impl.cuh"

// scratch of paths_trace_closest_static (below): held by every GPU that loads the library, shared by all calls
static __device__ __align__(256) unsigned char g_static_scratch[PATHS_STATIC_SCRATCH_BYTES];

namespace paths {
RendererBase *make_renderer_f32(const SceneDevice *sd, const paths_camera &cam, const paths_render_opts &opts, std::string &err) {
    return make_renderer_impl<float>(sd, cam, opts, err);
}
paths_status trace_closest_f32(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                               int kernel, std::string &err) {
    return trace_closest_impl<float>(sd, o, d, n, prim, t, bary, kernel, err);
}
paths_status trace_occluded_f32(const SceneDevice *sd, const double *o, const double *d, const uint32_t *light, size_t n, uint8_t *occluded,
                                int kernel, std::string &err) {
    return trace_occluded_impl<float>(sd, o, d, light, n, occluded, kernel, err);
}
// ------------------------------------------------------------------ paths_trace_closest_static: the per-ray seam without any allocation
// No counterpart in the reference; added for callers that may not call cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid it).  The scratch trace_closest_impl takes from cudaMalloc per call -- the
// staged rays, the path-state ray and hit records, the queue counters and the unstaged answers -- lives in ONE file-scope device array
// (g_static_scratch, top of this file) instead: every GPU that loads the library holds PATHS_STATIC_SCRATCH_BYTES (8.25 MiB) of it until exit.  All calls share the array
// and it keeps the last call's contents: every word a call reads is written by that call first (the rays and counters by copies from
// the host, a record per queue position by the kernels), so nothing has to start at zero except the fetch cursor, which is rewritten
// with the queue length every call.  Two calls that overlap on one GPU would race on it: calls are serialised by a lock here and each
// one waits for the device before it returns.

namespace {
std::mutex g_static_mu;
constexpr int kStaticMaxDevices = 64;
unsigned char *g_static_ptr[kStaticMaxDevices]; // device address of g_static_scratch on each ordinal, resolved once with that device current

// In host code `g_static_scratch` is the host-side stand-in of the symbol, not an address a kernel can use: ask the runtime.
paths_status static_scratch_address(int device, unsigned char **out, std::string &err) {
    if (device < 0 || device >= kStaticMaxDevices) {
        err = "device ordinal beyond the static scratch table";
        return PATHS_ERR_INVALID_ARG;
    }
    if (!g_static_ptr[device]) {
        PATHS_CU(cudaSetDevice(device));
        void *p = nullptr;
        PATHS_CU(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p || ((uintptr_t)p & 255u)) {
            err = "cudaGetSymbolAddress returned an unusable address for the static scratch array";
            return PATHS_ERR_CUDA;
        }
        g_static_ptr[device] = (unsigned char *)p;
    }
    *out = g_static_ptr[device];
    return PATHS_OK;
}
} // namespace

paths_status trace_closest_static_f32(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                                      std::string &err) {
    if (n == 0) return PATHS_OK;
    if (n > PATHS_STATIC_MAX_RAYS) {
        err = "paths_trace_closest_static: " + std::to_string(n) + " rays exceed the static scratch array's maximum of " +
              std::to_string((size_t)PATHS_STATIC_MAX_RAYS) + " per call (split the batch, or use paths_trace_closest)";
        return PATHS_ERR_INVALID_ARG;
    }
    std::lock_guard<std::mutex> lock(g_static_mu);
    PATHS_CU(cudaSetDevice(sd->device));
    unsigned char *base = nullptr;
    paths_status st = static_scratch_address(sd->device, &base, err);
    if (st != PATHS_OK) return st;
    // carve-up for the maximum batch, every piece 256-byte aligned: counters | origins | directions | t | barycentrics | prim | rays | hits
    const size_t N = PATHS_STATIC_MAX_RAYS;
    uint32_t *ctl = (uint32_t *)base;
    double *d_o = (double *)(base + 256);
    double *d_d = d_o + 3 * N;
    double *d_t = d_d + 3 * N;
    double *d_b = d_t + N;
    uint32_t *d_p = (uint32_t *)(d_b + 3 * N);
    float4 *rays = (float4 *)(d_p + N);
    float4 *hits = rays + 2 * N;
    static_assert(256 + (size_t)PATHS_STATIC_MAX_RAYS * (24 + 24 + 8 + 24 + 4 + 32 + 16) <= PATHS_STATIC_SCRATCH_BYTES, "static scratch too small");
    TraceConfig tc;
    st = configure_trace<float>(sd, tc, err); // occupancy queries and function attributes: no allocation
    if (st != PATHS_OK) return st;
    if (tc.variant == 16 && tc.grid4 == 0) {
        err = "the four-wide tree of this scene is not staged (or too deep for its shared-memory stack)";
        return PATHS_ERR_UNSUPPORTED;
    }
    const uint32_t host_ctl[2] = {(uint32_t)n, 0u}; // queue length, fetch cursor (the one word that must start at zero)
    PATHS_CU(cudaMemcpy(ctl, host_ctl, sizeof(host_ctl), cudaMemcpyHostToDevice));
    PATHS_CU(cudaMemcpy(d_o, o, n * 24, cudaMemcpyHostToDevice));
    PATHS_CU(cudaMemcpy(d_d, d, n * 24, cudaMemcpyHostToDevice));
    const unsigned blocks = (unsigned)((n + 127) / 128);
    k_stage_rays<float4><<<blocks, 128>>>(d_o, d_d, (uint32_t)n, rays);
    if (tc.variant == 16) PATHS_K_EXTEND4<<<tc.grid4, 128, tc.smem4>>>(sd->view, rays, hits, ctl, ctl + 1);
    else if (tc.variant == 3) PATHS_K_EXTEND2<<<tc.grid_trace, 128>>>(sd->view, rays, hits, ctl, ctl + 1);
    else k_extend<float, false><<<tc.grid_trace, 128>>>(sd->view, rays, hits, ctl, ctl + 1);
    k_unstage_hits<float4><<<blocks, 128>>>(hits, sd->view, (uint32_t)n, d_p, d_t, bary ? d_b : nullptr);
    PATHS_CU(cudaGetLastError());
    PATHS_CU(cudaDeviceSynchronize());
    PATHS_CU(cudaMemcpy(prim, d_p, n * 4, cudaMemcpyDeviceToHost));
    PATHS_CU(cudaMemcpy(t, d_t, n * 8, cudaMemcpyDeviceToHost));
    if (bary) PATHS_CU(cudaMemcpy(bary, d_b, n * 24, cudaMemcpyDeviceToHost));
    return PATHS_OK;
}

paths_status stack_overflows_f32(int device, unsigned long long *out, std::string &err) {
    PATHS_CU(cudaSetDevice(device));
    unsigned long long asserts = 0;
    return read_stack_overflows(out, &asserts, err);
}
paths_status camera_rays_f32(const paths_camera &cam, int device, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h,
                             double *o, double *d, double *weight, std::string &err) {
    return camera_rays_impl<float>(cam, device, seed, sample, x0, y0, w, h, o, d, weight, err);
}
} // namespace paths
