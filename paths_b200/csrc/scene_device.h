// scene_device.h -- a scene staged into one GPU's HBM (read-only, shared by every renderer on that device).
#pragma once
#include <string>
#include <vector>

#include "dev_scene.h"
#include "host_scene.h"

namespace paths {

struct SceneDevice {
    int device = -1;
    DevScene view;
    std::vector<void *> allocs;
    size_t bytes = 0;
    ~SceneDevice();
};

// Uploads the scene; ref / fast may be null (the matching kernels then must not be used).  The f32 mirrors of the shading tables
// go with the production tree (fast != null, or want_f32_tables when that tree is about to be built on the device).
SceneDevice *upload_scene(int device, const HostScene &hs, const RefTree *ref, const FastTree *fast, std::string &err,
                          bool want_f32_tables = false);

// ---- production tree built on the device (bvh_device.cu)
struct DeviceTreeInfo {
    float ms = 0.0f;                 // CUDA-event time of the whole build, scratch allocation included
    uint32_t nodes2 = 0, nodes4 = 0; // node slots allocated (two-wide: Karras index space, not all referenced)
    uint32_t depth2 = 0, depth4 = 0;
    uint32_t giants = 0, leaves = 0, hidden = 0;
};
// false + err == "small": too few ordinary primitives for a morton build, use the host builder
bool build_device_tree(SceneDevice *sd, const HostScene &hs, DeviceTreeInfo &info, std::string &err);
bool download_fast_tree(const SceneDevice *sd, const DeviceTreeInfo &info, std::vector<float> &nodes2, std::vector<float> &nodes4,
                        std::vector<float> &leaves, std::string &err);

struct RendererBase {
    virtual ~RendererBase() {}
    virtual paths_status render(uint32_t spp, bool wait) = 0;
    virtual paths_status sync() = 0;
    virtual paths_status image_f32(float *rgb) = 0;
    virtual paths_status image_rgb24(uint8_t *rgb) = 0;
    virtual paths_status get_sums(double *sums4) = 0;
    virtual paths_status add_sums(const double *sums4) = 0;
    virtual paths_status device_sums(void **ptr, size_t *n, int32_t *elem_bytes) = 0;
    virtual paths_status cuda_stream(void **stream) = 0;
    virtual paths_status get_stats(paths_stats *st) = 0;
    virtual paths_status set_camera(const double loc[3], const double rot[9]) = 0;
    virtual paths_status reset() = 0;
    virtual paths_status render_preview() = 0;
    virtual paths_status render_columns(uint32_t n_requests) = 0;
    virtual paths_status get_progress(uint32_t *samples_done, uint32_t *sweep_pos) = 0;
    virtual paths_status set_progress(uint32_t samples_done) = 0;
    // frame combine over NCCL (renderer.rs:93-105 is where the reference merges its workers' results): comm = ncclComm_t of this
    // renderer's rank; owned = destroy it with the renderer
    virtual paths_status attach_comm(void *comm, bool owned) = 0;
    virtual paths_status combine(int32_t root, bool wait) = 0;
    // the two halves of combine() for a caller that issues several ranks' NCCL calls inside one ncclGroupStart / End (paths_group):
    // NCCL defers grouped calls to the outermost GroupEnd, so what must run AFTER the transfer (the root's scatter) is launched after it
    virtual paths_status combine_post(int32_t root) = 0;   // NCCL calls (and the kernels that feed them)
    virtual paths_status combine_finish(int32_t root) = 0; // kernels that consume what was received
    virtual bool has_comm() const = 0;
    std::string err;
};

DevCamera make_dev_camera(const paths_camera &c); // CameraDescription::to_camera (src/serde.rs:172-188)

// one factory + two batch entry points per arithmetic type, instantiated in kernels_f32.cu / kernels_f64.cu
RendererBase *make_renderer_f32(const SceneDevice *sd, const paths_camera &cam, const paths_render_opts &opts, std::string &err);
RendererBase *make_renderer_f64(const SceneDevice *sd, const paths_camera &cam, const paths_render_opts &opts, std::string &err);
// kernel: 0 = the bounce-ray kernel a render launches for this scene, or a PATHS_TRACE_* code
paths_status trace_closest_f32(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                               int kernel, std::string &err);
// paths_trace_closest_static: the same kernels with all scratch in a file-scope device array of kernels_f32.cu (no allocation)
paths_status trace_closest_static_f32(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                                      std::string &err);
paths_status trace_occluded_f32(const SceneDevice *sd, const double *o, const double *d, const uint32_t *light, size_t n, uint8_t *occluded,
                                int kernel, std::string &err);
paths_status trace_occluded_f64(const SceneDevice *sd, const double *o, const double *d, const uint32_t *light, size_t n, uint8_t *occluded,
                                std::string &err);
paths_status trace_occluded_mixed(const SceneDevice *sd, const double *o, const double *d, const uint32_t *light, size_t n, uint8_t *occluded,
                                  std::string &err);
paths_status stack_overflows_f32(int device, unsigned long long *out, std::string &err);
paths_status stack_overflows_f64(int device, unsigned long long *out, std::string &err);
paths_status trace_closest_f64(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                               std::string &err);
paths_status trace_closest_mixed(const SceneDevice *sd, const double *o, const double *d, size_t n, uint32_t *prim, double *t, double *bary,
                                 std::string &err);
paths_status camera_rays_f32(const paths_camera &cam, int device, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h,
                             double *o, double *d, double *weight, std::string &err);
paths_status camera_rays_f64(const paths_camera &cam, int device, uint64_t seed, uint32_t sample, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h,
                             double *o, double *d, double *weight, std::string &err);

} // namespace paths
