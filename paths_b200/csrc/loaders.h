// loaders.h -- drop-in scene input: the YAML schema of src/serde.rs:65-385, Wavefront OBJ (src/obj.rs:8-67 via tobj 2.0.0),
// PLY (src/ply.rs:11-78 via ply-rs 0.1.2) and Radiance HDR (src/serde.rs:367-382 via image 0.23.4).  The three crates are
// third-party and not vendored in the reference; their published behaviour is restated (f32 coordinate parsing, fan
// triangulation, per-mesh vertex re-indexing, RGBE -> f32 as mantissa * 2^(e-136)).
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/paths_b200.h"

namespace paths {

struct LoadedModel {
    std::vector<double> vertices; // 3 per vertex (f32-rounded)
    std::vector<uint32_t> faces;  // 3 per face
    std::vector<double> colours;  // 3 per vertex or empty
    bool has_material = false;    // OBJ: MTL diffuse -> Lambertian (src/obj.rs:24-31)
    double kd[3] = {0, 0, 0};
};

struct LoadedScene {
    paths_camera camera;
    std::vector<LoadedModel> models;
    std::vector<paths_object_desc> objects;
    std::vector<paths_light_desc> lights;
    paths_sky_desc sky;
    std::vector<float> sky_texels;
    std::vector<paths_model_desc> model_descs; // filled by fill_desc
    void fill_desc(paths_scene_desc &d);
};

// ---- generic YAML subset (block + flow collections, scalars, comments, quotes)
struct YamlNode {
    enum Kind { Null, Scalar, Map, Seq } kind = Null;
    std::string scalar;
    bool quoted = false;
    std::vector<std::pair<std::string, std::shared_ptr<YamlNode>>> map; // insertion order kept
    std::vector<std::shared_ptr<YamlNode>> seq;
    int line = 0;
    const YamlNode *get(const std::string &key) const;
};
bool parse_yaml(const std::string &text, YamlNode &root, std::string &err);

paths_status load_scene_yaml(const std::string &yaml_path, const std::string &asset_root, LoadedScene &out, std::string &err);
paths_status load_obj(const std::string &path, std::vector<LoadedModel> &out, std::string &err);
paths_status load_ply(const std::string &path, LoadedModel &out, std::string &err);
paths_status load_hdr(const std::string &path, uint32_t &w, uint32_t &h, std::vector<float> &texels, std::string &err);

} // namespace paths
