// scene_yaml.cpp -- SceneDescription (src/serde.rs:65-385) read from YAML into the flattened C description.
//
// Accepted superset: the reference's own stale scenes (spheres_on_plane.yml, bokeh_demo.yml) omit
// `lights:`, the `type: Rgb` tag on albedo and Gloss `metalness`; they load here with lights = [], untagged albedo = Rgb,
// metalness = 0.
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <sys/stat.h>

#include "loaders.h"

namespace paths {

void LoadedScene::fill_desc(paths_scene_desc &d) {
    model_descs.resize(models.size());
    for (size_t i = 0; i < models.size(); i++) {
        model_descs[i].n_vertices = (uint32_t)(models[i].vertices.size() / 3);
        model_descs[i].n_faces = (uint32_t)(models[i].faces.size() / 3);
        model_descs[i].vertices = models[i].vertices.data();
        model_descs[i].faces = models[i].faces.data();
        model_descs[i].vertex_colours = models[i].colours.empty() ? nullptr : models[i].colours.data();
    }
    std::memset(&d, 0, sizeof(d));
    d.n_models = (uint32_t)models.size();
    d.n_objects = (uint32_t)objects.size();
    d.n_lights = (uint32_t)lights.size();
    d.models = model_descs.data();
    d.objects = objects.data();
    d.lights = lights.data();
    d.sky = sky;
    d.sky.texels = sky_texels.empty() ? nullptr : sky_texels.data();
}

namespace {

struct Ctx {
    std::string err;
    bool ok = true;
    void fail(const YamlNode *n, const std::string &msg) {
        if (ok) err = "scene yaml" + (n ? " line " + std::to_string(n->line) : std::string()) + ": " + msg;
        ok = false;
    }
    const YamlNode *req(const YamlNode *m, const char *key) {
        const YamlNode *c = m ? m->get(key) : nullptr;
        if (!c) fail(m, std::string("missing field `") + key + "`");
        return c;
    }
    double num(const YamlNode *n) {
        if (!n || n->kind != YamlNode::Scalar) {
            fail(n, "expected a number");
            return 0.0;
        }
        const std::string &s = n->scalar;
        if (s == ".inf" || s == "+.inf") return HUGE_VAL;
        if (s == "-.inf") return -HUGE_VAL;
        if (s == ".nan") return NAN;
        char *e = nullptr;
        double v = strtod(s.c_str(), &e);
        if (e == s.c_str() || *e != 0) fail(n, "expected a number, got '" + s + "'");
        return v;
    }
    double fnum(const YamlNode *m, const char *key) { return num(req(m, key)); }
    bool boolean(const YamlNode *n) {
        if (n && n->kind == YamlNode::Scalar) {
            if (n->scalar == "true" || n->scalar == "True") return true;
            if (n->scalar == "false" || n->scalar == "False") return false;
        }
        fail(n, "expected a boolean");
        return false;
    }
    std::string str(const YamlNode *n) {
        if (!n || n->kind != YamlNode::Scalar) {
            fail(n, "expected a string");
            return "";
        }
        return n->scalar;
    }
    void vec(const YamlNode *n, double out[3]) { // VectorDescription
        out[0] = fnum(n, "x");
        out[1] = fnum(n, "y");
        out[2] = fnum(n, "z");
    }
    void col(const YamlNode *n, double out[3]) { // ColourDescription
        out[0] = fnum(n, "r");
        out[1] = fnum(n, "g");
        out[2] = fnum(n, "b");
    }
    std::string tag(const YamlNode *n) { return str(req(n, "type")); }

    void albedo(const YamlNode *n, paths_basic_material &m) { // MaterialColourDescription (serde.rs:42-57)
        if (!n || n->kind != YamlNode::Map) {
            fail(n, "expected an albedo mapping");
            return;
        }
        const YamlNode *t = n->get("type");
        std::string ty = t ? str(t) : "Rgb";
        if (ty == "Vertex") m.albedo_mode = 1;
        else if (ty == "Rgb") col(n, m.albedo);
        else fail(n, "unknown variant `" + ty + "`, expected `Rgb` or `Vertex`");
    }
    void basic(const YamlNode *n, paths_basic_material &m) { // BasicMaterialDescription (serde.rs:265-271)
        std::memset(&m, 0, sizeof(m));
        std::string ty = tag(n);
        if (ty == "Lambertian") {
            m.type = PATHS_MAT_LAMBERTIAN;
            albedo(req(n, "albedo"), m);
        } else if (ty == "Gloss") {
            m.type = PATHS_MAT_GLOSS;
            albedo(req(n, "albedo"), m);
            m.reflectance = fnum(n, "reflectance");
            const YamlNode *mt = n->get("metalness");
            m.metalness = mt ? num(mt) : 0.0;
        } else if (ty == "Mirror") {
            m.type = PATHS_MAT_MIRROR;
        } else if (ty == "CookTorrance") {
            m.type = PATHS_MAT_COOK_TORRANCE;
            col(req(n, "albedo"), m.albedo);
            m.roughness = fnum(n, "roughness");
        } else fail(n, "unknown material type `" + ty + "`");
    }
    // returns false for `Auto`
    bool material(const YamlNode *n, paths_material_desc &m) { // MaterialDescription (serde.rs:254-263)
        std::memset(&m, 0, sizeof(m));
        m.refractive_index = 1.0;
        std::string ty = tag(n);
        if (ty == "Auto") return false;
        if (ty == "Fresnel") {
            m.m.type = PATHS_MAT_FRESNEL;
            m.refractive_index = fnum(n, "refractive_index");
            basic(req(n, "diffuse"), m.diffuse);
            basic(req(n, "specular"), m.specular);
        } else basic(n, m.m);
        return true;
    }
};

bool file_exists(const std::string &p) {
    struct stat st;
    return stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode);
}
std::string resolve_asset(const std::string &p, const std::string &root) {
    if (!p.empty() && p[0] == '/') return p;
    if (file_exists(p)) return p; // the reference opens paths relative to the working directory
    if (!root.empty() && file_exists(root + "/" + p)) return root + "/" + p;
    return p;
}
std::string lower_ext(const std::string &p) {
    size_t d = p.find_last_of('.');
    std::string e = d == std::string::npos ? "" : p.substr(d + 1);
    for (char &c : e) c = (char)tolower((unsigned char)c);
    return e;
}

} // namespace

paths_status load_scene_yaml(const std::string &yaml_path, const std::string &asset_root_in, LoadedScene &out, std::string &err) {
    std::ifstream f(yaml_path, std::ios::binary);
    if (!f) {
        err = "cannot open " + yaml_path;
        return PATHS_ERR_IO;
    }
    std::ostringstream ss;
    ss << f.rdbuf();
    YamlNode root;
    if (!parse_yaml(ss.str(), root, err)) return PATHS_ERR_PARSE;
    std::string asset_root = asset_root_in;
    if (asset_root.empty()) { // default: the directory above the YAML's directory (scenes/x.yml -> repo root)
        size_t s1 = yaml_path.find_last_of('/');
        std::string dir = s1 == std::string::npos ? "." : yaml_path.substr(0, s1);
        asset_root = dir + "/..";
    }
    Ctx c;
    out = LoadedScene();
    if (root.kind != YamlNode::Map) {
        err = "scene yaml: top level is not a mapping";
        return PATHS_ERR_PARSE;
    }
    // ---- camera (serde.rs:156-188)
    const YamlNode *cam = c.req(&root, "camera");
    std::memset(&out.camera, 0, sizeof(out.camera));
    if (cam) {
        {
            const double iw = c.fnum(cam, "image_width"), ih = c.fnum(cam, "image_height");
            const bool in_range = iw >= 1.0 && iw <= 65536.0 && ih >= 1.0 && ih <= 65536.0; // false for NaN too
            if (!in_range) c.fail(cam, "image_width / image_height out of range");
            out.camera.image_width = in_range ? (uint32_t)iw : 0u; // the cast is only taken where it is defined
            out.camera.image_height = in_range ? (uint32_t)ih : 0u;
        }
        c.vec(c.req(cam, "location"), out.camera.location);
        const YamlNode *o = c.req(cam, "orientation");
        out.camera.pitch = c.fnum(o, "pitch");
        out.camera.yaw = c.fnum(o, "yaw");
        out.camera.roll = c.fnum(o, "roll");
        out.camera.sensor_width = c.fnum(cam, "sensor_width");
        out.camera.sensor_height = c.fnum(cam, "sensor_height");
        out.camera.focal_length = c.fnum(cam, "focal_length");
        out.camera.focus_distance = c.fnum(cam, "focus_distance");
        out.camera.aperture = c.fnum(cam, "aperture");
    }
    if (!c.ok) {
        err = c.err;
        return PATHS_ERR_PARSE;
    }
    // ---- models: name -> file (serde.rs:72-73,87-90); loaded on first use by an object (model.rs:49-94)
    std::map<std::string, std::string> declared;
    if (const YamlNode *models = root.get("models")) {
        if (models->kind == YamlNode::Map)
            for (const auto &kv : models->map) declared[kv.first] = c.str(c.req(kv.second.get(), "file"));
        else if (models->kind != YamlNode::Null) c.fail(models, "`models` must be a mapping");
    }
    std::map<std::string, std::vector<int>> loaded;
    // ---- objects (serde.rs:91-142)
    const YamlNode *objects = c.req(&root, "objects");
    if (objects && objects->kind == YamlNode::Seq) {
        for (const auto &on : objects->seq) {
            const YamlNode *shape = c.req(on.get(), "shape");
            const YamlNode *matn = c.req(on.get(), "material");
            if (!c.ok) break;
            std::string st = c.tag(shape);
            paths_object_desc od;
            std::memset(&od, 0, sizeof(od));
            if (st == "Sphere") {
                od.shape = 0;
                c.vec(c.req(shape, "center"), od.center);
                od.radius = c.fnum(shape, "radius");
                od.scale = 1.0;
                if (!c.material(matn, od.material) && c.ok) c.fail(matn, "Cannot directly convert Auto material description into material."); // serde.rs:276
                out.objects.push_back(od);
            } else if (st == "Mesh") {
                std::string name = c.str(c.req(shape, "model"));
                if (!c.ok) break;
                if (!declared.count(name)) {
                    c.fail(shape, "Attempt to load model '" + name + "' before declaration"); // model.rs:55-58
                    break;
                }
                if (!loaded.count(name)) {
                    std::string fp = resolve_asset(declared[name], asset_root);
                    std::string ext = lower_ext(fp);
                    std::vector<LoadedModel> ms;
                    paths_status s;
                    if (ext == "obj") s = load_obj(fp, ms, err);
                    else if (ext == "ply") {
                        ms.resize(1);
                        s = load_ply(fp, ms[0], err);
                    } else {
                        err = "Unknown file extension: " + ext; // model.rs:80
                        return PATHS_ERR_UNSUPPORTED;
                    }
                    if (s != PATHS_OK) return s;
                    std::vector<int> ids;
                    for (LoadedModel &m : ms) {
                        ids.push_back((int)out.models.size());
                        out.models.push_back(std::move(m));
                    }
                    loaded[name] = ids;
                }
                od.shape = 1;
                const YamlNode *sm = shape->get("smooth_normals");
                od.smooth_normals = sm ? (c.boolean(sm) ? 1 : 0) : 1; // serde.rs:243-252
                c.vec(c.req(shape, "translation"), od.translation);
                const YamlNode *rot = c.req(shape, "rotation");
                od.rot_pitch = c.fnum(rot, "pitch");
                od.rot_yaw = c.fnum(rot, "yaw");
                od.rot_roll = c.fnum(rot, "roll");
                od.scale = c.fnum(shape, "scale");
                paths_material_desc md;
                bool explicit_mat = c.material(matn, md);
                for (int ix : loaded[name]) { // one Object per sub-model (serde.rs:112-141)
                    od.model = ix;
                    if (explicit_mat) od.material = md;
                    else { // Auto: the model's own material, else white Lambertian (serde.rs:127-132)
                        std::memset(&od.material, 0, sizeof(od.material));
                        od.material.refractive_index = 1.0;
                        od.material.m.type = PATHS_MAT_LAMBERTIAN;
                        const LoadedModel &lm = out.models[ix];
                        for (int k = 0; k < 3; k++) od.material.m.albedo[k] = lm.has_material ? lm.kd[k] : 1.0;
                    }
                    out.objects.push_back(od);
                }
            } else c.fail(shape, "unknown variant `" + st + "`, expected `Sphere` or `Mesh`");
            if (!c.ok) break;
        }
    } else if (objects) c.fail(objects, "`objects` must be a sequence");
    // ---- lights (serde.rs:144-151, 203-229); absent in the stale scenes
    if (const YamlNode *lights = root.get("lights")) {
        if (lights->kind == YamlNode::Seq) {
            for (const auto &ln : lights->seq) {
                paths_light_desc ld;
                std::memset(&ld, 0, sizeof(ld));
                const YamlNode *g = c.req(ln.get(), "geometry");
                if (!c.ok) break;
                std::string gt = c.tag(g);
                if (gt == "Point") {
                    ld.geometry = 0;
                    c.vec(g, ld.v);
                } else if (gt == "Sphere") {
                    ld.geometry = 1;
                    c.vec(c.req(g, "center"), ld.v);
                    ld.radius = c.fnum(g, "radius");
                } else c.fail(g, "unknown variant `" + gt + "`, expected `Point` or `Sphere`");
                c.col(c.req(ln.get(), "colour"), ld.colour);
                ld.intensity = c.fnum(ln.get(), "intensity");
                out.lights.push_back(ld);
            }
        } else if (lights->kind != YamlNode::Null) c.fail(lights, "`lights` must be a sequence");
    }
    // ---- skybox (serde.rs:336-384)
    const YamlNode *sky = c.req(&root, "skybox");
    std::memset(&out.sky, 0, sizeof(out.sky));
    if (sky && c.ok) {
        std::string t = c.tag(sky);
        if (t == "Flat") {
            out.sky.type = 0;
            c.col(c.req(sky, "colour"), out.sky.c0);
        } else if (t == "Gradient") {
            out.sky.type = 1;
            c.col(c.req(sky, "overhead_colour"), out.sky.c0);
            c.col(c.req(sky, "horizon_colour"), out.sky.c1);
        } else if (t == "Hdri") {
            out.sky.type = 2;
            std::string fp = resolve_asset(c.str(c.req(sky, "filename")), asset_root);
            if (c.ok) {
                paths_status s = load_hdr(fp, out.sky.width, out.sky.height, out.sky_texels, err);
                if (s != PATHS_OK) return s;
            }
        } else c.fail(sky, "unknown variant `" + t + "`, expected one of `Flat`, `Gradient`, `Hdri`");
    }
    if (!c.ok) {
        err = c.err;
        return PATHS_ERR_PARSE;
    }
    return PATHS_OK;
}

} // namespace paths
