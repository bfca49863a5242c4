// kernels_f32.cu -- production instantiation of the wavefront pipeline (f32 traversal and shading, f64 spheres).
#include "renderer_impl.cuh"

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
