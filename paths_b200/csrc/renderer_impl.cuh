// renderer_impl.cuh -- RendererImpl<R>: buffers, pass loop and estimator for one GPU (see renderer.cuh).
#pragma once
#include <cstdlib>

#include "renderer.cuh"

#ifndef PATHS_MIXED_REFILL
#define PATHS_MIXED_REFILL 1 /* 0 = plain batches (k_extend / k_shadow), 1 = lane refill for bounce and visibility rays, 2 = camera rays too */
#endif
#ifndef PATHS_SHARD_TILE
#define PATHS_SHARD_TILE 32u /* edge of the square image tiles dealt to the ranks under PATHS_SHARD_TILES */
#endif
#ifndef PATHS_SAMPLE_GROUP_LOG2
/* samples per batch of 32 slots = 2^this (ShardInfo::split).  Same box, whole frame, M samples/s (profiles/r02q_sg_ab.txt): teapot 3945 / 3987 / 4042 /
 * 4047, environment.yml 7477 / 7677 / 7575 / 7742, dragon 1132 / 1137 / 1143 / 1150, bunny 2884 / 2882 / 2907 / 2946 for 1 / 2 / 4 / 8 samples */
#define PATHS_SAMPLE_GROUP_LOG2 3u
#endif
#ifndef PATHS_EXTEND2_KI
#define PATHS_EXTEND2_KI 12
#endif

namespace paths {

template <typename R> class RendererImpl : public RendererBase {
    typedef typename Vec4T<R>::type R4;

    const SceneDevice *sd_;
    paths_render_opts opts_;
    DevCamera cam_;
    cudaStream_t stream_ = nullptr; // the renderer's stream: everything a caller can observe is ordered on it
    // Two pass contexts on two streams: consecutive passes of one render() call overlap, so that the thin late bounces of one pass
    // (and the drain of every persistent kernel) run next to the wide early bounces of the other.  Accumulation stays in pass order.
    struct PassCtx {
        PathState<R4> ps;
        PassControl pc;
        uint32_t *ctl = nullptr;
        cudaStream_t stream = nullptr;
        cudaEvent_t acc = nullptr; // recorded after this context's last k_accumulate
        bool acc_valid = false;
    };
    PassCtx ctx_[4];
    int n_ctx_ = 1, cur_ = 0, next_ctx_ = 0;
    cudaStream_t cs_ = nullptr; // stream of the context in use
    cudaEvent_t fork_ = nullptr, join_ = nullptr;
    void use_ctx(int i) {
        cur_ = i;
        ps_ = ctx_[i].ps;
        pc_ = ctx_[i].pc;
        ctl_block_ = ctx_[i].ctl;
        cs_ = ctx_[i].stream;
    }
    uint32_t W_, H_, npix_, n_local_ = 0, spp_pass_ = 1, n_ctl_ = 0;
    size_t n_slots_max_ = 0;
    bool mixed_ = false;        // f64 instantiation over the production tree (PATHS_PRECISION_MIXED)
    uint32_t samples_done_ = 0; // global samples per pixel issued so far (all ranks together)
    TraceConfig tc_; // which traversal kernels this scene runs on (shared with the per-ray seam, renderer.cuh)
    int grid_shade_ = 0, grid_stream_ = 0, shade_variant_ = 1;
    bool camera_tiles_ = true;
    uint32_t sample_group_log2_ = PATHS_SAMPLE_GROUP_LOG2;
    // device buffers
    R4 *sums_ = nullptr;
    // frame combine (NCCL): the communicator of this rank, the combined estimator on the root, tile-shard staging
    ncclComm_t comm_ = nullptr;
    bool comm_owned_ = false;
    R4 *frame_ = nullptr;        // [npix_] combined estimator, allocated on the first combine this rank is the root of
    bool frame_valid_ = false;   // the image / estimator reads show frame_ (set by combine on the root, cleared by any new sample)
    R4 *pack_ = nullptr;         // tile shards: this rank's owned pixels, dense (send side) / every other rank's (root side)
    size_t pack_elems_ = 0;
    uint32_t *peer_pixels_ = nullptr; // root, tile shards: pixel lists of the other ranks, concatenated in rank order
    int peer_pixels_root_ = -1;
    std::vector<uint32_t> peer_counts_;
    size_t scatter_total_ = 0;
    uint32_t *pixels_ = nullptr; // tile shard: local -> global pixel
    // interactive front half (renderer.rs:143-192): column sweeps from the centre outwards, preview grid
    uint32_t *sweep_pixels_ = nullptr; // scratch local -> global table for column / preview passes [npix_]
    uint32_t sweep_pos_ = 0;           // Renderer::block_num: requests issued in the current sweep
    uint32_t n_previews_ = 0;
    PathState<R4> ps_;
    PassControl pc_;
    uint32_t *ctl_block_ = nullptr;
    void *stage_ = nullptr; // staging buffer of the image / estimator transfers
    size_t stage_bytes_ = 0;
    std::vector<void *> allocs_;
    // stats
    paths_stats stats_;
    struct Timed {
        cudaEvent_t a, b;
        int cls;
    };
    std::vector<Timed> timed_;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> spans_; // one per render() call
    std::vector<cudaEvent_t> pool_;

    template <typename T> paths_status alloc(T **p, size_t count) {
        void *q = nullptr;
        PATHS_CU(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
        allocs_.push_back(q);
        *p = (T *)q;
        return PATHS_OK;
    }
    cudaEvent_t get_event() {
        if (!pool_.empty()) {
            cudaEvent_t e = pool_.back();
            pool_.pop_back();
            return e;
        }
        cudaEvent_t e;
        cudaEventCreate(&e);
        return e;
    }
    void tic(int cls) {
        if (!opts_.time_kernels) return;
        Timed t;
        t.a = get_event();
        t.b = get_event();
        t.cls = cls;
        cudaEventRecord(t.a, cs_);
        timed_.push_back(t);
    }
    void toc() {
        stats_.kernel_launches++;
        if (!opts_.time_kernels) return;
        cudaEventRecord(timed_.back().b, cs_);
    }
    void resolve_timers() {
        for (Timed &t : timed_) {
            float ms = 0.0f;
            cudaEventElapsedTime(&ms, t.a, t.b);
            double *dst[] = {&stats_.ms_raygen, &stats_.ms_extend, &stats_.ms_shade, &stats_.ms_shadow, &stats_.ms_accumulate};
            *dst[t.cls] += ms;
            stats_.ms_total += ms;
            pool_.push_back(t.a);
            pool_.push_back(t.b);
        }
        timed_.clear();
        for (auto &sp : spans_) {
            float ms = 0.0f;
            cudaEventElapsedTime(&ms, sp.first, sp.second);
            stats_.ms_device += ms;
            pool_.push_back(sp.first);
            pool_.push_back(sp.second);
        }
        spans_.clear();
    }

  public:
    RendererImpl(const SceneDevice *sd, const paths_camera &cam, const paths_render_opts &o) : sd_(sd), opts_(o) {
        cam_ = make_dev_camera(cam);
        mixed_ = IsF64<R>::value && o.precision == PATHS_PRECISION_MIXED;
        W_ = cam.image_width;
        H_ = cam.image_height;
        npix_ = W_ * H_;
        std::memset(&stats_, 0, sizeof(stats_));
        std::memset(&ps_, 0, sizeof(ps_));
        std::memset(&pc_, 0, sizeof(pc_));
    }
    ~RendererImpl() override {
        cudaSetDevice(sd_->device);
        if (stream_) cudaStreamSynchronize(stream_);
        for (Timed &t : timed_) {
            cudaEventDestroy(t.a);
            cudaEventDestroy(t.b);
        }
        for (auto &sp : spans_) {
            cudaEventDestroy(sp.first);
            cudaEventDestroy(sp.second);
        }
        for (cudaEvent_t e : pool_) cudaEventDestroy(e);
        if (comm_ && comm_owned_) {
            std::string e;
            if (const NcclApi *api = nccl_api(e)) api->CommDestroy(comm_);
        }
        for (void *p : allocs_) cudaFree(p);
        if (stage_) cudaFree(stage_);
        for (int c = 1; c < 4; c++)
            if (ctx_[c].stream) {
                cudaStreamSynchronize(ctx_[c].stream);
                cudaStreamDestroy(ctx_[c].stream);
            }
        for (PassCtx &c : ctx_)
            if (c.acc) cudaEventDestroy(c.acc);
        if (fork_) cudaEventDestroy(fork_);
        if (join_) cudaEventDestroy(join_);
        if (stream_) cudaStreamDestroy(stream_);
    }

    paths_status init() {
        PATHS_CU(cudaSetDevice(sd_->device));
        PATHS_CU(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
        uint32_t world = std::max<uint32_t>(opts_.world, 1);
        if (opts_.rank >= world) {
            err = "rank >= world";
            return PATHS_ERR_INVALID_ARG;
        }
        // ---- shard: which pixels this rank owns
        if (opts_.shard_mode == PATHS_SHARD_TILES && world > 1) {
            std::vector<uint32_t> own;
            owned_pixels(W_, H_, opts_.rank, world, own);
            n_local_ = (uint32_t)own.size();
            if (alloc(&pixels_, own.size()) != PATHS_OK) return PATHS_ERR_CUDA;
            if (!own.empty()) PATHS_CU(cudaMemcpy(pixels_, own.data(), own.size() * 4, cudaMemcpyHostToDevice));
        } else {
            n_local_ = npix_;
        }
        // ---- wavefront size: about 32M paths in flight unless told otherwise (4.5 GB of path state in f32: late bounces of a
        // pass run on few rays, and bigger passes amortise those tails -- measured +15% from 8M to 32M slots at 1080p)
        spp_pass_ = opts_.samples_per_pass;
        if (spp_pass_ == 0) spp_pass_ = (uint32_t)std::min<uint64_t>(64, std::max<uint64_t>(1, (32u << 20) / std::max<uint32_t>(n_local_, 1)));
        n_slots_max_ = (size_t)std::max<uint32_t>(n_local_, 1) * spp_pass_;
        if (n_slots_max_ > 0x7FFFFFFFull) {
            err = "samples_per_pass * pixels exceeds 2^31 slots";
            return PATHS_ERR_INVALID_ARG;
        }
        if (alloc(&sums_, npix_) != PATHS_OK) return PATHS_ERR_CUDA;
        PATHS_CU(cudaMemsetAsync(sums_, 0, (size_t)npix_ * sizeof(R4), stream_));
        // pass contexts: the second one only when kernels are not timed one by one (overlap would blur per-kernel times)
        n_ctx_ = opts_.time_kernels ? 1 : 2;
#ifdef PATHS_DEV_SWITCHES
        if (const char *v = getenv("PATHS_PIPELINE")) n_ctx_ = opts_.time_kernels ? 1 : std::min(4, std::max(1, atoi(v)));
#endif
        unsigned long long *totals = nullptr;
        if (alloc(&totals, 4) != PATHS_OK) return PATHS_ERR_CUDA;
        PATHS_CU(cudaMemsetAsync(totals, 0, 4 * sizeof(unsigned long long), stream_));
        const uint32_t nb = opts_.max_bounces + 2;
        n_ctl_ = (3 + PATHS_NWORK) * nb;
        for (int c = 0; c < n_ctx_; c++) {
            PassCtx &x = ctx_[c];
            std::memset(&x.ps, 0, sizeof(x.ps));
            std::memset(&x.pc, 0, sizeof(x.pc));
            // records dense by queue position, two parities (dev_scene.h): {origin | slot, direction | flags} pairs and throughputs
            for (int p = 0; p < 2; p++)
                if (alloc(&x.ps.ray[p], 2 * n_slots_max_) != PATHS_OK || alloc(&x.ps.thr[p], n_slots_max_) != PATHS_OK) return PATHS_ERR_CUDA;
            if (alloc(&x.ps.hit, n_slots_max_) != PATHS_OK || alloc(&x.ps.col, n_slots_max_) != PATHS_OK) return PATHS_ERR_CUDA;
            // shadow records: one per path and bounce for light NEE, one more with environment NEE
            size_t n_shadow = n_slots_max_ * ((opts_.extensions & PATHS_EXT_ENV_NEE) ? 2 : 1);
            R4 **shadow[] = {&x.ps.sh_o, &x.ps.sh_d, &x.ps.sh_c};
            for (R4 **p : shadow)
                if (alloc(p, n_shadow) != PATHS_OK) return PATHS_ERR_CUDA;
            if (alloc(&x.ctl, n_ctl_) != PATHS_OK) return PATHS_ERR_CUDA;
            x.pc.qcount = x.ctl;
            x.pc.scount = x.ctl + nb;
            x.pc.ecount = x.ctl + 2 * nb;
            x.pc.work = x.ctl + 3 * nb;
            x.pc.totals = totals;
            if (c == 0) x.stream = stream_;
            else PATHS_CU(cudaStreamCreateWithFlags(&x.stream, cudaStreamNonBlocking));
            PATHS_CU(cudaEventCreateWithFlags(&x.acc, cudaEventDisableTiming));
        }
        PATHS_CU(cudaEventCreateWithFlags(&fork_, cudaEventDisableTiming));
        PATHS_CU(cudaEventCreateWithFlags(&join_, cudaEventDisableTiming));
        use_ctx(0);
        // ---- traversal kernels and persistent grids (every SM full): the selection the per-ray seam shares
        {
            paths_status st = configure_trace<R>(sd_, tc_, err);
            if (st != PATHS_OK) return st;
        }
#ifdef PATHS_DEV_SWITCHES
        if (const char *v = getenv("PATHS_SHADE_VARIANT")) shade_variant_ = atoi(v);
        if (const char *v = getenv("PATHS_CAMERA_TILES")) camera_tiles_ = atoi(v) != 0;
        if (const char *v = getenv("PATHS_SAMPLE_GROUP_LOG2")) sample_group_log2_ = (uint32_t)std::min(5, std::max(0, atoi(v)));
#endif
        int bps = 0;
        {
            const bool lights = sd_->view.n_lights > 0, ext = opts_.extensions != 0;
            if (ext) PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (k_shade<R, true, true>), 128, 0));
            else if (lights) PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (k_shade<R, true, false>), 128, 0));
            else PATHS_CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (k_shade<R, false, false>), 128, 0));
        }
        grid_shade_ = tc_.sm_count * std::max(bps, 1);
        grid_stream_ = tc_.sm_count * 8;
        return PATHS_OK;
    }

    void launch_extend(uint32_t b) {
        const R4 *q = ps_.ray[b & 1u]; // the dense ray records this bounce reads
        if constexpr (!IsF64<R>::value) {
            const uint32_t *cnt = pc_.qcount + b;
            uint32_t *cur = pc_.work + PATHS_NWORK * b;
            int vsel = (b == 0 && tc_.variant_b0 >= 0) ? tc_.variant_b0 : tc_.variant;
            switch (vsel) { // 0 = one batch of 32 rays per warp at a time (k_extend), kept as the plain baseline
#ifdef PATHS_EXTEND2_SPEC
            case 3: k_extend2s<1, 16, PATHS_EXTEND2_KI, PATHS_EXTEND2_SPEC><<<tc_.grid_trace, 128, 0, cs_>>>(sd_->view, q, ps_.hit, cnt, cur); return;
#else
            case 3: PATHS_K_EXTEND2<<<tc_.grid_trace, 128, 0, cs_>>>(sd_->view, q, ps_.hit, cnt, cur); return;
#endif
            case 16: PATHS_K_EXTEND4<<<tc_.grid4, 128, tc_.smem4, cs_>>>(sd_->view, q, ps_.hit, cnt, cur); return;
            default: break;
            }
        }
        if constexpr (IsF64<R>::value) {
            // mixed mode: lane refill for the bounce rays; the camera segment keeps plain batches (coherent rays, like k_camera_extend)
            if (mixed_ && PATHS_MIXED_REFILL && (b > 0 || PATHS_MIXED_REFILL > 1)) {
                if (tc_.grid_mixed4 > 0)
                    PATHS_K_EXTEND4M<<<tc_.grid_mixed4, 128, tc_.smem4, cs_>>>(sd_->view, q, ps_.hit, pc_.qcount + b, pc_.work + PATHS_NWORK * b);
                else PATHS_K_EXTEND2M<<<tc_.grid_mixed, 128, 0, cs_>>>(sd_->view, q, ps_.hit, pc_.qcount + b, pc_.work + PATHS_NWORK * b);
                return;
            }
        }
        if (mixed_) k_extend<R, true><<<tc_.grid_trace, 128, 0, cs_>>>(sd_->view, q, ps_.hit, pc_.qcount + b, pc_.work + PATHS_NWORK * b);
        else k_extend<R, false><<<tc_.grid_trace, 128, 0, cs_>>>(sd_->view, q, ps_.hit, pc_.qcount + b, pc_.work + PATHS_NWORK * b);
    }
    // visibility of the records [0, *count) of `ps`'s shadow arrays; `cursor` is this launch's dynamic-fetch counter
    void launch_shadow(const PathState<R4> &ps, const uint32_t *count, uint32_t *cursor) {
        if constexpr (!IsF64<R>::value) {
            if (tc_.variant == 16) {
                PATHS_K_SHADOW4<<<tc_.grid4, 128, tc_.smem4, cs_>>>(sd_->view, ps, count, cursor, pc_.totals);
                return;
            }
            if (tc_.variant != 0) {
                PATHS_K_SHADOW2<<<tc_.grid_trace, 128, 0, cs_>>>(sd_->view, ps, count, cursor, pc_.totals);
                return;
            }
        }
        if constexpr (IsF64<R>::value) {
            if (mixed_ && PATHS_MIXED_REFILL) {
                if (tc_.grid_mixed4 > 0) PATHS_K_SHADOW4M<<<tc_.grid_mixed4, 128, tc_.smem4, cs_>>>(sd_->view, ps, count, cursor, pc_.totals);
                else PATHS_K_SHADOW2M<<<tc_.grid_mixed, 128, 0, cs_>>>(sd_->view, ps, count, cursor, pc_.totals);
                return;
            }
        }
        if (mixed_) k_shadow<R, true><<<tc_.grid_trace, 128, 0, cs_>>>(sd_->view, ps, count, cursor, pc_.totals);
        else k_shadow<R, false><<<tc_.grid_trace, 128, 0, cs_>>>(sd_->view, ps, count, cursor, pc_.totals);
    }

    paths_status run_pass(uint32_t s_base, uint32_t n_spp) { return run_pass_on(pixels_, n_local_, s_base, n_spp); }

    // one wavefront pass over `n_local` pixels (`pixels` = local -> global index, or null for the whole frame in order)
    paths_status run_pass_on(const uint32_t *pixels, uint32_t n_local, uint32_t s_base, uint32_t n_spp) {
        const uint32_t n_slots = n_local * n_spp;
        if (n_slots == 0) return PATHS_OK;
        frame_valid_ = false;
        const uint32_t k0 = (uint32_t)opts_.seed, k1 = (uint32_t)(opts_.seed >> 32);
        ShardInfo shard;
        shard.n_local = n_local;
        shard.pixels = pixels;
        shard.n_local_magic = n_local > 1 ? (uint32_t)((1ull << 32) / n_local) : 0xFFFFFFFFu;
        // whole frame: local pixels numbered in 8 x 4 tile order when the size allows it (ShardInfo::pixel_of)
        shard.tiles_x = 0;
        shard.tiles_x_magic = 0;
        if (camera_tiles_ && pixels == nullptr && n_local == npix_ && W_ % 8 == 0 && H_ % 4 == 0 && npix_ >= 32 &&
            (uint64_t)(npix_ / 32) * (W_ / 8) < (1ull << 32)) { // the multiply-high division by tiles_x is exact below that
            shard.tiles_x = W_ / 8;
            shard.tiles_x_magic = (uint32_t)((1ull << 32) / shard.tiles_x) + 1u;
        }
        // sample groups: 2^g samples of 32 >> g neighbouring pixels per batch of 32 slots (ShardInfo)
        shard.sg_log2 = 0;
        shard.n_blocks = 0;
        shard.n_blocks_magic = 0;
        uint32_t g = sample_group_log2_;
        while (g > 0 && n_spp % (1u << g) != 0) g--; // the largest group size that divides this pass's sample count
        if (g > 0 && n_local % 32 == 0) {
            shard.sg_log2 = g;
            shard.n_blocks = n_local >> (5u - shard.sg_log2);
            shard.n_blocks_magic = shard.n_blocks > 1 ? (uint32_t)((1ull << 32) / shard.n_blocks) : 0xFFFFFFFFu;
        }
        const uint32_t nb = opts_.max_bounces + 1; // segments per path (trace.rs:13-16)
        PATHS_CU(cudaMemsetAsync(ctl_block_, 0, n_ctl_ * sizeof(uint32_t), cs_));
        bool fused_camera = false;
        if constexpr (!IsF64<R>::value) fused_camera = tc_.variant_b0 == 1;
        if (!fused_camera) {
            tic(0);
            k_raygen<R><<<grid_stream_, 256, 0, cs_>>>(cam_, k0, k1, shard, s_base, n_slots, ps_, pc_);
            toc();
        }
        for (uint32_t b = 0; b < nb; b++) {
            tic(1);
            if (b == 0 && fused_camera) {
                if constexpr (!IsF64<R>::value) {
                    if (tc_.camera_wide4 && tc_.grid_camera4 > 0)
                        k_camera_extend<float, true><<<tc_.grid_camera4, 128, tc_.smem4, cs_>>>(sd_->view, cam_, k0, k1, shard, s_base, n_slots, ps_, pc_, pc_.work);
                    else
                        k_camera_extend<float, false><<<tc_.grid_camera, 128, 0, cs_>>>(sd_->view, cam_, k0, k1, shard, s_base, n_slots, ps_, pc_, pc_.work);
                }
            } else launch_extend(b);
            toc();
            stats_.n_extend_launches++;
            tic(2);
            {
                uint32_t *shade_cursor = shade_variant_ == 1 ? nullptr : pc_.work + PATHS_NWORK * b + 1;
                // three instantiations: every extension compiled in (lights or not) / lights only / neither
#define PATHS_SHADE_ARGS sd_->view, ps_, ps_.ray[b & 1u], ps_.thr[b & 1u], ps_.ray[(b + 1u) & 1u], ps_.thr[(b + 1u) & 1u], shard, s_base, k0, k1, opts_.max_bounces, opts_.extensions, b, pc_.qcount + b, pc_.qcount + b + 1, pc_.scount + b, pc_.ecount + b, (uint32_t)n_slots_max_, shade_cursor, pc_.totals
                if (opts_.extensions != 0) {
                    if (sd_->view.n_lights > 0) k_shade<R, true, true><<<grid_shade_, 128, 0, cs_>>>(PATHS_SHADE_ARGS);
                    else k_shade<R, false, true><<<grid_shade_, 128, 0, cs_>>>(PATHS_SHADE_ARGS);
                } else if (sd_->view.n_lights > 0) k_shade<R, true, false><<<grid_shade_, 128, 0, cs_>>>(PATHS_SHADE_ARGS);
                else k_shade<R, false, false><<<grid_shade_, 128, 0, cs_>>>(PATHS_SHADE_ARGS);
#undef PATHS_SHADE_ARGS
            }
            toc();
            if (sd_->view.n_lights > 0) {
                tic(3);
                launch_shadow(ps_, pc_.scount + b, pc_.work + PATHS_NWORK * b + 2);
                toc();
                stats_.n_shadow_launches++;
            }
            // environment NEE records (extension) live in the second half of the shadow arrays and are resolved by a launch of their
            // own: a path may emit a light record AND an environment record at one bounce, and two records of one slot inside one
            // launch would race on its radiance record (plain load-add-store).  Stream order makes the sum deterministic.
            if ((opts_.extensions & PATHS_EXT_ENV_NEE) && sd_->view.sky_type == 2) {
                PathState<R4> pe = ps_;
                pe.sh_o += n_slots_max_;
                pe.sh_d += n_slots_max_;
                pe.sh_c += n_slots_max_;
                tic(3);
                launch_shadow(pe, pc_.ecount + b, pc_.work + PATHS_NWORK * b + 3);
                toc();
                stats_.n_shadow_launches++;
            }
        }
        // the estimator is shared by the pass contexts: accumulate in pass order (the previous pass ran on the previous context)
        if (n_ctx_ > 1 && ctx_[(cur_ + n_ctx_ - 1) % n_ctx_].acc_valid)
            PATHS_CU(cudaStreamWaitEvent(cs_, ctx_[(cur_ + n_ctx_ - 1) % n_ctx_].acc, 0));
        tic(4);
        k_accumulate<R><<<grid_stream_, 256, 0, cs_>>>(ps_, shard, n_spp, sums_);
        toc();
        if (n_ctx_ > 1) {
            PATHS_CU(cudaEventRecord(ctx_[cur_].acc, cs_));
            ctx_[cur_].acc_valid = true;
        }
        k_tally<R><<<1, 32, 0, cs_>>>(pc_, nb, n_slots);
        stats_.kernel_launches++;
        stats_.passes++;
        PATHS_CU(cudaGetLastError());
        return PATHS_OK;
    }

    paths_status render(uint32_t spp, bool wait) override {
        PATHS_CU(cudaSetDevice(sd_->device));
        if (sweep_pos_ != 0) {
            err = "a column sweep is in progress: finish it with render_columns() or reset() before render()";
            return PATHS_ERR_INVALID_ARG;
        }
        uint32_t world = std::max<uint32_t>(opts_.world, 1);
        uint32_t s0 = samples_done_, s1 = samples_done_ + spp;
        if (opts_.shard_mode == PATHS_SHARD_SAMPLES && world > 1) { // this rank's slice of the sample range
            uint32_t lo = (uint32_t)((uint64_t)spp * opts_.rank / world), hi = (uint32_t)((uint64_t)spp * (opts_.rank + 1) / world);
            s0 = samples_done_ + lo;
            s1 = samples_done_ + hi;
        }
        cudaEvent_t ea = get_event(), eb = get_event();
        cudaEventRecord(ea, stream_);
        const bool overlap = n_ctx_ > 1 && s1 - s0 > spp_pass_;
        if (overlap) { // work already queued on the renderer's stream (a reset, a caller's copy) comes first on both streams
            PATHS_CU(cudaEventRecord(fork_, stream_));
            for (int c = 1; c < n_ctx_; c++) PATHS_CU(cudaStreamWaitEvent(ctx_[c].stream, fork_, 0));
        }
        // passes go round the contexts; the round continues across calls so that accumulation order stays pass order
        for (uint32_t s = s0; s < s1; s += spp_pass_) {
            use_ctx(overlap ? next_ctx_ : 0);
            if (overlap) next_ctx_ = (next_ctx_ + 1) % n_ctx_;
            paths_status st = run_pass(s, std::min(spp_pass_, s1 - s));
            if (st != PATHS_OK) {
                use_ctx(0);
                return st;
            }
        }
        use_ctx(0);
        if (overlap) { // join: whatever follows on the renderer's stream sees every pass
            for (int c = 1; c < n_ctx_; c++) {
                PATHS_CU(cudaEventRecord(join_, ctx_[c].stream));
                PATHS_CU(cudaStreamWaitEvent(stream_, join_, 0));
            }
        }
        cudaEventRecord(eb, stream_);
        spans_.push_back(std::make_pair(ea, eb));
        samples_done_ += spp;
        return wait ? sync() : PATHS_OK;
    }
    paths_status sync() override {
        PATHS_CU(cudaSetDevice(sd_->device));
        PATHS_CU(cudaStreamSynchronize(stream_));
        resolve_timers();
        return PATHS_OK;
    }
    // one staging buffer for every image / estimator transfer, grown on demand (no cudaMalloc / cudaFree per frame: both
    // synchronise the whole device, which costs tens of milliseconds once NCCL communicators are alive)
    paths_status staging(size_t bytes, void **out) {
        if (bytes > stage_bytes_) {
            if (stage_) cudaFree(stage_);
            stage_ = nullptr;
            stage_bytes_ = 0;
            PATHS_CU(cudaMalloc(&stage_, bytes));
            stage_bytes_ = bytes;
        }
        *out = stage_;
        return PATHS_OK;
    }
    paths_status image_f32(float *rgb) override {
        PATHS_CU(cudaSetDevice(sd_->device));
        float *d = nullptr;
        if (staging((size_t)npix_ * 12, (void **)&d) != PATHS_OK) return PATHS_ERR_CUDA;
        k_image_f32<R><<<(npix_ + 255) / 256, 256, 0, stream_>>>(shown(), npix_, W_, d);
        PATHS_CU(cudaMemcpyAsync(rgb, d, (size_t)npix_ * 12, cudaMemcpyDeviceToHost, stream_));
        PATHS_CU(cudaStreamSynchronize(stream_));
        return PATHS_OK;
    }
    paths_status image_rgb24(uint8_t *rgb) override {
        PATHS_CU(cudaSetDevice(sd_->device));
        uint8_t *d = nullptr;
        if (staging((size_t)npix_ * 3, (void **)&d) != PATHS_OK) return PATHS_ERR_CUDA;
        k_image_rgb24<R><<<(npix_ + 255) / 256, 256, 0, stream_>>>(shown(), npix_, W_, d);
        PATHS_CU(cudaMemcpyAsync(rgb, d, (size_t)npix_ * 3, cudaMemcpyDeviceToHost, stream_));
        PATHS_CU(cudaStreamSynchronize(stream_));
        return PATHS_OK;
    }
    paths_status get_sums(double *sums4) override {
        PATHS_CU(cudaSetDevice(sd_->device));
        double4 *d = nullptr;
        if (staging((size_t)npix_ * 32, (void **)&d) != PATHS_OK) return PATHS_ERR_CUDA;
        k_sums_to_f64<R><<<(npix_ + 255) / 256, 256, 0, stream_>>>(shown(), npix_, d);
        PATHS_CU(cudaMemcpyAsync(sums4, d, (size_t)npix_ * 32, cudaMemcpyDeviceToHost, stream_));
        PATHS_CU(cudaStreamSynchronize(stream_));
        return PATHS_OK;
    }
    paths_status add_sums(const double *sums4) override {
        PATHS_CU(cudaSetDevice(sd_->device));
        double4 *d = nullptr;
        if (staging((size_t)npix_ * 32, (void **)&d) != PATHS_OK) return PATHS_ERR_CUDA;
        frame_valid_ = false;
        PATHS_CU(cudaMemcpyAsync(d, sums4, (size_t)npix_ * 32, cudaMemcpyHostToDevice, stream_));
        k_add_sums<R><<<(npix_ + 255) / 256, 256, 0, stream_>>>(sums_, npix_, d);
        PATHS_CU(cudaStreamSynchronize(stream_));
        return PATHS_OK;
    }
    // what Renderer::render shows: the combined frame right after a combine on its root, this rank's own estimator otherwise
    const R4 *shown() const { return frame_valid_ ? frame_ : sums_; }

    // ---- frame combine ----------------------------------------------------------------------------------------------------
    bool has_comm() const override { return comm_ != nullptr; }
    paths_status attach_comm(void *comm, bool owned) override {
        if (comm_) {
            err = "a communicator is already attached to this renderer";
            return PATHS_ERR_INVALID_ARG;
        }
        comm_ = (ncclComm_t)comm;
        comm_owned_ = owned;
        return PATHS_OK;
    }
    // The pixels of `rank` under tile sharding: PATHS_SHARD_TILE x PATHS_SHARD_TILE tiles (32: at 64 the 64 tiles a rank of 8 gets at
    // 1080p left its time 6.5 % above the mean on environment.yml -- sky tiles against mesh tiles; profiles/r02n_bench_n8_environment_tiles.json)
    // dealt round-robin in row-major tile order (tile t -> rank t mod world).  Listed tile by tile, and inside a tile in 8 x 4 blocks, so
    // that 32 consecutive local pixels are a compact block of the image (same reason as ShardInfo::pixel_of).  Root and senders derive the
    // same list, so the combine needs no exchange of indices.
    static void owned_pixels(uint32_t W, uint32_t H, uint32_t rank, uint32_t world, std::vector<uint32_t> &own) {
        own.clear();
        const uint32_t T = PATHS_SHARD_TILE, tiles_x = (W + T - 1) / T, tiles_y = (H + T - 1) / T;
        for (uint32_t ty = 0; ty < tiles_y; ty++)
            for (uint32_t tx = 0; tx < tiles_x; tx++) {
                if ((ty * tiles_x + tx) % world != rank) continue;
                for (uint32_t by = 0; by < T; by += 4)
                    for (uint32_t bx = 0; bx < T; bx += 8)
                        for (uint32_t in = 0; in < 32; in++) { // Morton order inside the block, like ShardInfo::pixel_of
                            const uint32_t ix = (in & 1u) | ((in >> 1) & 2u) | ((in >> 2) & 4u), iy = ((in >> 1) & 1u) | ((in >> 2) & 2u);
                            const uint32_t x = tx * T + bx + ix, y = ty * T + by + iy;
                            if (x < W && y < H) own.push_back(x + y * W);
                        }
            }
    }
    // Renderer::drain_result_queue -> Estimator::update_pixel (renderer.rs:93-105) across GPUs: partial estimators of all ranks are
    // combined on `root`, enqueued on the renderer's stream behind its passes.  Sample shards: ncclReduce(sum) of the {sum r, g, b, n}
    // buffers.  Tile shards: every rank owns disjoint pixels, so the owners' values are MOVED, not added: each rank packs its pixels into
    // a dense buffer and sends it, the root receives the others' buffers in one NCCL group and scatters them -- bit-identical to one GPU
    // and 1/world of the bytes a reduce of full frames moves.
    paths_status combine(int32_t root, bool wait) override {
        paths_status st = combine_post(root);
        if (st == PATHS_OK) st = combine_finish(root);
        if (st != PATHS_OK) return st;
        return wait ? sync() : PATHS_OK;
    }
    paths_status combine_finish(int32_t root) override {
        const uint32_t world = std::max<uint32_t>(opts_.world, 1);
        if (world == 1 || !comm_) return PATHS_OK;
        PATHS_CU(cudaSetDevice(sd_->device));
        const bool is_root = (uint32_t)root == opts_.rank;
        if (is_root && opts_.shard_mode == PATHS_SHARD_TILES && scatter_total_ > 0) {
            k_scatter_pixels<R4><<<(unsigned)((scatter_total_ + 255) / 256), 256, 0, stream_>>>(pack_, peer_pixels_, (uint32_t)scatter_total_, frame_);
            stats_.kernel_launches++;
        }
        PATHS_CU(cudaGetLastError());
        frame_valid_ = is_root;
        return PATHS_OK;
    }
    paths_status combine_post(int32_t root) override {
        const uint32_t world = std::max<uint32_t>(opts_.world, 1);
        scatter_total_ = 0;
        if (world == 1) return PATHS_OK; // nothing to combine: the estimator is the frame
        if (!comm_) {
            err = "no communicator attached: call paths_renderer_attach_comm (or render through a paths_group)";
            return PATHS_ERR_INVALID_ARG;
        }
        if (root < 0 || (uint32_t)root >= world) {
            err = "combine root out of range";
            return PATHS_ERR_INVALID_ARG;
        }
        std::string e;
        const NcclApi *api = nccl_api(e);
        if (!api) {
            err = e;
            return PATHS_ERR_NCCL;
        }
        PATHS_CU(cudaSetDevice(sd_->device));
        const bool is_root = (uint32_t)root == opts_.rank;
        const ncclDataType_t dt = IsF64<R>::value ? ncclDouble : ncclFloat;
        if (is_root && !frame_ && alloc(&frame_, npix_) != PATHS_OK) return PATHS_ERR_CUDA;
        if (opts_.shard_mode == PATHS_SHARD_SAMPLES) {
            PATHS_NCCL(api, Reduce(sums_, is_root ? frame_ : nullptr, (size_t)npix_ * 4, dt, ncclSum, root, comm_, stream_));
        } else if (!is_root) {
            if (pack_elems_ < n_local_) {
                if (alloc(&pack_, n_local_) != PATHS_OK) return PATHS_ERR_CUDA;
                pack_elems_ = n_local_;
            }
            if (n_local_ > 0) {
                k_pack_pixels<R4><<<(n_local_ + 255) / 256, 256, 0, stream_>>>(sums_, pixels_, n_local_, pack_);
                stats_.kernel_launches++;
                PATHS_NCCL(api, Send(pack_, (size_t)n_local_ * 4, dt, root, comm_, stream_));
            }
        } else {
            if (peer_pixels_root_ != root) { // the other ranks' pixel lists follow from the tile dealing: no exchange needed
                std::vector<uint32_t> all, own;
                peer_counts_.assign(world, 0);
                for (uint32_t r = 0; r < world; r++) {
                    if (r == opts_.rank) continue;
                    owned_pixels(W_, H_, r, world, own);
                    peer_counts_[r] = (uint32_t)own.size();
                    all.insert(all.end(), own.begin(), own.end());
                }
                if (alloc(&peer_pixels_, all.size()) != PATHS_OK) return PATHS_ERR_CUDA;
                if (!all.empty()) PATHS_CU(cudaMemcpy(peer_pixels_, all.data(), all.size() * 4, cudaMemcpyHostToDevice));
                if (pack_elems_ < all.size()) {
                    if (alloc(&pack_, all.size()) != PATHS_OK) return PATHS_ERR_CUDA;
                    pack_elems_ = all.size();
                }
                peer_pixels_root_ = root;
            }
            // own pixels (and zeros elsewhere) first, then the others' pixels over them
            PATHS_CU(cudaMemcpyAsync(frame_, sums_, (size_t)npix_ * sizeof(R4), cudaMemcpyDeviceToDevice, stream_));
            size_t off = 0, total = 0;
            PATHS_NCCL(api, GroupStart());
            for (uint32_t r = 0; r < world; r++) {
                if (r == opts_.rank || peer_counts_[r] == 0) continue;
                ncclResult_t rr = api->Recv(pack_ + off, (size_t)peer_counts_[r] * 4, dt, (int)r, comm_, stream_);
                if (rr != ncclSuccess) {
                    api->GroupEnd();
                    err = std::string("ncclRecv: ") + api->GetErrorString(rr);
                    return PATHS_ERR_NCCL;
                }
                off += peer_counts_[r];
            }
            PATHS_NCCL(api, GroupEnd());
            total = off;
            scatter_total_ = total; // scattered into frame_ by combine_finish, once the receives are really enqueued
        }
        PATHS_CU(cudaGetLastError());
        return PATHS_OK;
    }

    paths_status device_sums(void **ptr, size_t *n, int32_t *elem_bytes) override {
        *ptr = sums_;
        *n = (size_t)npix_ * 4;
        *elem_bytes = (int32_t)sizeof(R);
        return PATHS_OK;
    }
    paths_status cuda_stream(void **stream) override {
        *stream = (void *)stream_;
        return PATHS_OK;
    }
    paths_status get_stats(paths_stats *st) override {
        PATHS_CU(cudaSetDevice(sd_->device));
        PATHS_CU(cudaStreamSynchronize(stream_));
        resolve_timers();
        unsigned long long t[4];
        PATHS_CU(cudaMemcpy(t, pc_.totals, sizeof(t), cudaMemcpyDeviceToHost));
        stats_.extend_rays = t[0];
        stats_.shadow_rays = t[1];
        stats_.bad_samples = t[2];
        stats_.samples = t[3];
        {
            unsigned long long ovf = 0, asserts = 0;
            paths_status so = read_stack_overflows(&ovf, &asserts, err);
            if (so != PATHS_OK) return so;
            stats_.stack_overflows = ovf;
            stats_.assert_failures = asserts;
#ifdef PATHS_DEBUG_ASSERTS
            stats_.debug_asserts = 1;
#endif
        }
        stats_.extend_kernel = tc_.variant;
        stats_.camera_kernel = tc_.variant_b0;
        stats_.shade_kernel = shade_variant_;
        stats_.pass_contexts = n_ctx_;
        *st = stats_;
        return PATHS_OK;
    }
    paths_status set_camera(const double loc[3], const double rot[9]) override { // renderer.rs:112-120
        for (int i = 0; i < 3; i++) cam_.location[i] = loc[i];
        for (int i = 0; i < 9; i++) cam_.rot[i] = rot[i];
        return reset();
    }
    paths_status reset() override { // renderer.rs:122-128,143-150: a new epoch drops the estimator
        PATHS_CU(cudaSetDevice(sd_->device));
        PATHS_CU(cudaStreamSynchronize(stream_));
        PATHS_CU(cudaMemsetAsync(sums_, 0, (size_t)npix_ * sizeof(R4), stream_));
        PATHS_CU(cudaMemsetAsync(pc_.totals, 0, 4 * sizeof(unsigned long long), stream_));
        samples_done_ = 0;
        sweep_pos_ = 0;
        n_previews_ = 0;
        frame_valid_ = false;
        resolve_timers();
        std::memset(&stats_, 0, sizeof(stats_));
        return PATHS_OK;
    }

    // ---- interactive front half -------------------------------------------------------------------------------------------
    paths_status ensure_sweep_table() {
        if (!sweep_pixels_ && alloc(&sweep_pixels_, npix_) != PATHS_OK) return PATHS_ERR_CUDA;
        return PATHS_OK;
    }
    paths_status single_gpu_only(const char *what) {
        if (std::max<uint32_t>(opts_.world, 1) > 1) {
            err = std::string(what) + " is a single-GPU call (world must be 1)";
            return PATHS_ERR_UNSUPPORTED;
        }
        return PATHS_OK;
    }
    // Renderer::request_preview (renderer.rs:152-164): one sample for every 6th pixel in x and y, accumulated like any other
    // sample; Estimator::render then fills pixels without samples from that grid (pixels.rs:53-79, see k_image_*).
    paths_status render_preview() override {
        PATHS_CU(cudaSetDevice(sd_->device));
        paths_status st = single_gpu_only("render_preview");
        if (st != PATHS_OK) return st;
        if ((st = ensure_sweep_table()) != PATHS_OK) return st;
        std::vector<uint32_t> px;
        for (uint32_t x = 0; x < W_; x += PATHS_PREVIEW_GRID) // column-major like the reference's request order
            for (uint32_t y = 0; y < H_; y += PATHS_PREVIEW_GRID) px.push_back(x + y * W_);
        PATHS_CU(cudaMemcpyAsync(sweep_pixels_, px.data(), px.size() * 4, cudaMemcpyHostToDevice, stream_));
        PATHS_CU(cudaStreamSynchronize(stream_)); // px is a stack-owned staging buffer
        // preview samples use their own index range so that they never repeat a sample of the regular schedule
        st = run_pass_on(sweep_pixels_, (uint32_t)px.size(), 0x40000000u + n_previews_, 1);
        n_previews_++;
        return st == PATHS_OK ? sync() : st;
    }
    // Renderer::next_request + fill/drain (renderer.rs:79-110,166-192): the next `n_requests` pixel columns, from the centre of the
    // image outwards; one sample per pixel ((1,1) pattern) during the first sweep of an epoch, 25 ((5,5) pattern) afterwards.
    // For odd widths the reference's formula asks for the centre column twice and never for column 0; here the repeat goes to
    // column 0, so that every column is rendered once per sweep.
    paths_status render_columns(uint32_t n_requests) override {
        PATHS_CU(cudaSetDevice(sd_->device));
        paths_status st = single_gpu_only("render_columns");
        if (st != PATHS_OK) return st;
        if ((st = ensure_sweep_table()) != PATHS_OK) return st;
        std::vector<uint32_t> px;
        while (n_requests > 0) {
            const uint32_t n_spp = samples_done_ == 0 ? 1u : 25u; // quick_render (renderer.rs:172-176)
            const uint32_t max_cols = std::max<uint32_t>(1, (uint32_t)std::min<size_t>(n_slots_max_ / ((size_t)H_ * n_spp), 0xFFFFFFFFu));
            if ((size_t)H_ * n_spp > n_slots_max_) {
                err = "samples_per_pass too small for one 25-sample column request";
                return PATHS_ERR_INVALID_ARG;
            }
            uint32_t take = std::min({n_requests, W_ - sweep_pos_, max_cols});
            px.clear();
            for (uint32_t k = 0; k < take; k++) {
                uint32_t n = sweep_pos_ + k;
                uint32_t x = (n % 2 == 0) ? (W_ + n) / 2 : (W_ - n) / 2; // renderer.rs:170
                if ((W_ & 1u) && n == 1) x = 0;                          // odd width: the repeated centre column (see above)
                for (uint32_t y = 0; y < H_; y++) px.push_back(x + y * W_);
            }
            PATHS_CU(cudaMemcpyAsync(sweep_pixels_, px.data(), px.size() * 4, cudaMemcpyHostToDevice, stream_));
            PATHS_CU(cudaStreamSynchronize(stream_));
            if ((st = run_pass_on(sweep_pixels_, (uint32_t)px.size(), samples_done_, n_spp)) != PATHS_OK) return st;
            PATHS_CU(cudaStreamSynchronize(stream_)); // the table is rewritten by the next chunk
            sweep_pos_ += take;
            n_requests -= take;
            if (sweep_pos_ >= W_) { // renderer.rs:180-183
                sweep_pos_ = 0;
                samples_done_ += n_spp;
            }
        }
        return sync();
    }
    paths_status get_progress(uint32_t *samples_done, uint32_t *sweep_pos) override {
        *samples_done = samples_done_;
        *sweep_pos = sweep_pos_;
        return PATHS_OK;
    }
    paths_status set_progress(uint32_t samples_done) override { // checkpoint resume: continue the sample schedule where it stopped
        samples_done_ = samples_done;
        sweep_pos_ = 0;
        return PATHS_OK;
    }
};

template <typename R>
RendererBase *make_renderer_impl(const SceneDevice *sd, const paths_camera &cam, const paths_render_opts &opts, std::string &err) {
    RendererImpl<R> *r = new RendererImpl<R>(sd, cam, opts);
    if (r->init() != PATHS_OK) {
        err = r->err;
        delete r;
        return nullptr;
    }
    return r;
}

} // namespace paths
