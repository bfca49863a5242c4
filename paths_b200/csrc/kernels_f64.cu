// kernels_f64.cu -- parity instantiation: f64, the reference's tree and operation order.  Compiled with -fmad=false
// (rustc never contracts a*b+c), so distances and primitive ids are bit-identical to the reference algorithm.
#include "renderer_impl.cuh"

namespace paths {
RendererBase *make_renderer_f64(const SceneDevice *sd, const paths_camera &cam, const paths_render_opts &opts, std::string &err) {
    return make_renderer_impl<double>(sd, cam, opts, err);
}
paths_status trace_closest_f64(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                               std::string &err) {
    return trace_closest_impl<double>(sd, o, d, n, prim, t, bary, 0, err);
}
paths_status trace_occluded_f64(const SceneDevice *sd, const double *o, const double *d, const uint32_t *light, size_t n, uint8_t *occluded,
                                std::string &err) {
    return trace_occluded_impl<double>(sd, o, d, light, n, occluded, 0, err);
}
paths_status trace_occluded_mixed(const SceneDevice *sd, const double *o, const double *d, const uint32_t *light, size_t n, uint8_t *occluded,
                                  std::string &err) {
    return trace_occluded_impl<double, true>(sd, o, d, light, n, occluded, 0, err);
}
paths_status stack_overflows_f64(int device, unsigned long long *out, std::string &err) {
    PATHS_CU(cudaSetDevice(device));
    unsigned long long asserts = 0;
    return read_stack_overflows(out, &asserts, err);
}
paths_status trace_closest_mixed(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                                 std::string &err) {
    return trace_closest_impl<double, true>(sd, o, d, n, prim, t, bary, 0, err);
}
paths_status camera_rays_f64(const paths_camera &cam, int device, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h,
                             double *o, double *d, double *weight, std::string &err) {
    return camera_rays_impl<double>(cam, device, seed, sample, x0, y0, w, h, o, d, weight, err);
}
} // namespace paths
