// cli_main.cpp -- headless stand-in for the reference's SDL2 front end (src/main.rs:39-186): scene YAML in, image out.
//   paths_cli scene.yml [--spp N] [--width W --height H] [--out img.ppm] [--pfm img.pfm] [--seed S] [--device D]
//             [--precision 32|64] [--rank R --world N --shard samples|tiles]
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
    if (argc < 2) {
        std::fprintf(stderr, "usage: %s scene.yml [--spp N] [--width W] [--height H] [--out img.ppm] [--pfm img.pfm] [--seed S] [--device D] "
                             "[--precision 32|64] [--rank R] [--world N] [--shard samples|tiles]\n", argv[0]);
        return 2;
    }
    std::string scene_path = argv[1], out = "out.ppm", pfm;
    uint32_t spp = 64, width = 0, height = 0;
    paths_render_opts opts;
    paths_render_opts_default(&opts);
    for (int i = 2; i + 1 < argc; i += 2) {
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
        else if (k == "--shard") opts.shard_mode = std::strcmp(v, "tiles") == 0 ? PATHS_SHARD_TILES : PATHS_SHARD_SAMPLES;
        else {
            std::fprintf(stderr, "unknown option %s\n", k.c_str());
            return 2;
        }
    }
    paths_scene *scene = nullptr;
    paths_camera cam;
    if (paths_scene_load_yaml(scene_path.c_str(), nullptr, width, height, &scene, &cam) != PATHS_OK) return die("load");
    paths_renderer *r = nullptr;
    if (paths_renderer_create(scene, &cam, &opts, &r) != PATHS_OK) return die("renderer");
    auto t0 = std::chrono::steady_clock::now();
    if (paths_renderer_render(r, spp) != PATHS_OK) return die("render");
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
