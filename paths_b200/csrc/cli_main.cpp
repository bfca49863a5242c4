// cli_main.cpp -- headless stand-in for the reference's SDL2 front end (src/main.rs:39-186): scene YAML in, image out.
//   paths_cli [scene.yml] [--spp N] [--width W --height H] [--out img.ppm] [--pfm img.pfm] [--seed S] [--device D]
//             [--precision 32|64] [--rank R --world N --shard samples|tiles] [--bvh host|device] [--stress N]
//             [--resume file.ck] [--checkpoint file.ck] [--sweeps K] [--gpus N]
// --gpus N renders one frame on devices 0 .. N-1 of this process (sample or tile shards) and combines the partial estimators on
// device 0 over NCCL inside the library (paths_group_*); the image is the combined frame.
// Without a scene file the seeded 500-sphere stress scene is rendered, like the reference does (src/main.rs:47-50).
// --sweeps K renders progressively instead of --spp: a preview pass, then K centre-out column sweeps (src/renderer.rs:152-192).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/paths_b200.h"

static int die(const char *what) {
    std::fprintf(stderr, "paths_cli: %s: %s\n", what, paths_last_error());
    return 1;
}

int main(int argc, char **argv) {
    if (argc >= 2 && (std::strcmp(argv[1], "-h") == 0 || std::strcmp(argv[1], "--help") == 0)) {
        std::fprintf(stderr, "usage: %s [scene.yml] [--spp N] [--width W] [--height H] [--out img.ppm] [--pfm img.pfm] [--seed S] [--device D] "
                             "[--precision 32|64] [--rank R] [--world N] [--shard samples|tiles] [--bvh host|device] [--stress N] "
                             "[--resume file.ck] [--checkpoint file.ck] [--sweeps K] [--gpus N]\n", argv[0]);
        return 2;
    }
    int first_opt = (argc >= 2 && std::strncmp(argv[1], "--", 2) != 0) ? 2 : 1;
    std::string scene_path = first_opt == 2 ? argv[1] : "", out = "out.ppm", pfm, resume, checkpoint;
    uint32_t spp = 64, width = 0, height = 0, stress = 500, sweeps = 0;
    int gpus = 0;
    bool device_bvh = false;
    paths_render_opts opts;
    paths_render_opts_default(&opts);
    for (int i = first_opt; i + 1 < argc; i += 2) {
        std::string k = argv[i];
        const char *v = argv[i + 1];
        if (k == "--spp") spp = (uint32_t)atoi(v);
        else if (k == "--width") width = (uint32_t)atoi(v);
        else if (k == "--height") height = (uint32_t)atoi(v);
        else if (k == "--out") out = v;
        else if (k == "--pfm") pfm = v;
        else if (k == "--seed") opts.seed = strtoull(v, nullptr, 10);
        else if (k == "--device") opts.device = atoi(v);
        else if (k == "--precision") opts.precision = atoi(v);
        else if (k == "--rank") opts.rank = (uint32_t)atoi(v);
        else if (k == "--world") opts.world = (uint32_t)atoi(v);
        else if (k == "--bvh") device_bvh = std::strcmp(v, "device") == 0;
        else if (k == "--stress") stress = (uint32_t)atoi(v);
        else if (k == "--resume") resume = v;
        else if (k == "--checkpoint") checkpoint = v;
        else if (k == "--sweeps") sweeps = (uint32_t)atoi(v);
        else if (k == "--gpus") gpus = atoi(v);
        else if (k == "--shard") opts.shard_mode = std::strcmp(v, "tiles") == 0 ? PATHS_SHARD_TILES : PATHS_SHARD_SAMPLES;
        else {
            std::fprintf(stderr, "unknown option %s\n", k.c_str());
            return 2;
        }
    }
    paths_scene *scene = nullptr;
    paths_camera cam;
    if (scene_path.empty()) {
        std::printf("No scene file passed in, generating random stress scene...\n"); // src/main.rs:48
        if (paths_scene_create_stress(stress, opts.seed, &scene, &cam) != PATHS_OK) return die("stress scene");
        if (width) cam.image_width = width;
        if (height) cam.image_height = height;
    } else if (paths_scene_load_yaml(scene_path.c_str(), nullptr, width, height, &scene, &cam) != PATHS_OK) return die("load");
    if (device_bvh && paths_scene_build_bvh(scene, PATHS_BVH_DEVICE) != PATHS_OK) return die("bvh");
    if (gpus > 1) { // one frame on several GPUs of this process: shards rendered per device, combined on device 0 inside the library
        if (sweeps > 0 || !resume.empty() || !checkpoint.empty()) {
            std::fprintf(stderr, "paths_cli: --gpus renders whole frames: --sweeps / --resume / --checkpoint are single-GPU options\n");
            return 2;
        }
        paths_group *g = nullptr;
        if (paths_group_create(scene, &cam, &opts, gpus, nullptr, &g) != PATHS_OK) return die("group");
        auto t0 = std::chrono::steady_clock::now();
        if (paths_group_render(g, spp) != PATHS_OK) return die("render");
        double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        paths_stats st;
        if (paths_group_get_stats(g, &st) != PATHS_OK) return die("stats");
        std::printf("Num rays: %llu (avg %.1f per pixel); %d GPUs; %.3f s; %.2f Msamples/s; %.2f Mrays/s\n", (unsigned long long)st.samples,
                    (double)st.samples / ((double)cam.image_width * cam.image_height), gpus, sec, st.samples / sec * 1e-6,
                    (st.extend_rays + st.shadow_rays) / sec * 1e-6);
        std::vector<uint8_t> rgb((size_t)cam.image_width * cam.image_height * 3);
        if (paths_group_image_rgb24(g, rgb.data()) != PATHS_OK) return die("image");
        if (paths_write_ppm(out.c_str(), rgb.data(), cam.image_width, cam.image_height) != PATHS_OK) return die("write");
        if (!pfm.empty()) {
            std::vector<float> f((size_t)cam.image_width * cam.image_height * 3);
            if (paths_group_image_f32(g, f.data()) != PATHS_OK) return die("image");
            if (paths_write_pfm(pfm.c_str(), f.data(), cam.image_width, cam.image_height) != PATHS_OK) return die("write");
        }
        paths_group_destroy(g);
        paths_scene_destroy(scene);
        return 0;
    }
    paths_renderer *r = nullptr;
    if (paths_renderer_create(scene, &cam, &opts, &r) != PATHS_OK) return die("renderer");
    if (!resume.empty() && paths_renderer_load_checkpoint(r, resume.c_str()) != PATHS_OK) return die("resume");
    auto t0 = std::chrono::steady_clock::now();
    if (sweeps > 0) {
        if (paths_renderer_render_preview(r) != PATHS_OK) return die("preview");
        if (paths_renderer_render_columns(r, sweeps * cam.image_width) != PATHS_OK) return die("render");
    } else if (paths_renderer_render(r, spp) != PATHS_OK) return die("render");
    if (!checkpoint.empty() && paths_renderer_save_checkpoint(r, checkpoint.c_str()) != PATHS_OK) return die("checkpoint");
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    paths_stats st;
    paths_renderer_get_stats(r, &st);
    // the reference prints "[elapsed][fps] Num rays: N (avg n per pixel)" (src/main.rs:111); N counts camera samples
    std::printf("Num rays: %llu (avg %.1f per pixel); %.3f s; %.2f Msamples/s; %.2f Mrays/s\n", (unsigned long long)st.samples,
                (double)st.samples / ((double)cam.image_width * cam.image_height), sec, st.samples / sec * 1e-6,
                (st.extend_rays + st.shadow_rays) / sec * 1e-6);
    std::vector<uint8_t> rgb((size_t)cam.image_width * cam.image_height * 3);
    if (paths_renderer_image_rgb24(r, rgb.data()) != PATHS_OK) return die("image");
    if (paths_write_ppm(out.c_str(), rgb.data(), cam.image_width, cam.image_height) != PATHS_OK) return die("write");
    if (!pfm.empty()) {
        std::vector<float> f((size_t)cam.image_width * cam.image_height * 3);
        if (paths_renderer_image_f32(r, f.data()) != PATHS_OK) return die("image");
        if (paths_write_pfm(pfm.c_str(), f.data(), cam.image_width, cam.image_height) != PATHS_OK) return die("write");
    }
    paths_renderer_destroy(r);
    paths_scene_destroy(scene);
    return 0;
}
