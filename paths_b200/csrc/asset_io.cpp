// asset_io.cpp -- OBJ / PLY / Radiance HDR readers (see loaders.h for what each restates).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

#include "loaders.h"

namespace paths {
namespace {

bool read_file(const std::string &path, std::string &out) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    std::ostringstream ss;
    ss << f.rdbuf();
    out = ss.str();
    return true;
}

std::vector<std::string> split_ws(const std::string &s) {
    std::vector<std::string> out;
    size_t i = 0;
    while (i < s.size()) {
        while (i < s.size() && isspace((unsigned char)s[i])) i++;
        size_t b = i;
        while (i < s.size() && !isspace((unsigned char)s[i])) i++;
        if (i > b) out.push_back(s.substr(b, i - b));
    }
    return out;
}

std::string dir_of(const std::string &path) {
    size_t p = path.find_last_of('/');
    return p == std::string::npos ? std::string(".") : path.substr(0, p);
}

std::string join_from(const std::vector<std::string> &tok, size_t from) {
    std::string s;
    for (size_t i = from; i < tok.size(); i++) {
        if (i > from) s += " ";
        s += tok[i];
    }
    return s;
}

} // namespace

// ------------------------------------------------------------------ OBJ (src/obj.rs:8-67; tobj::load_obj(file, triangulate = true))
paths_status load_obj(const std::string &path, std::vector<LoadedModel> &out, std::string &err) {
    std::string text;
    if (!read_file(path, text)) {
        err = "cannot open " + path;
        return PATHS_ERR_IO;
    }
    std::vector<float> pos; // f32, as tobj parses them
    std::vector<int64_t> cur_faces;
    std::map<std::string, std::vector<double>> mats;
    std::string cur_mat;
    bool have_mat = false;
    paths_status status = PATHS_OK;

    auto flush = [&]() {
        if (cur_faces.empty()) return;
        LoadedModel m;
        // tobj re-indexes the positions each mesh uses, in order of first use
        std::vector<int64_t> remap(pos.size() / 3, -1);
        for (int64_t v : cur_faces) {
            if (remap[v] < 0) {
                remap[v] = (int64_t)(m.vertices.size() / 3);
                m.vertices.push_back((double)pos[3 * v]);
                m.vertices.push_back((double)pos[3 * v + 1]);
                m.vertices.push_back((double)pos[3 * v + 2]);
            }
            m.faces.push_back((uint32_t)remap[v]);
        }
        if (have_mat && mats.count(cur_mat)) { // src/obj.rs:24-31: diffuse colour -> Lambertian
            m.has_material = true;
            for (int i = 0; i < 3; i++) m.kd[i] = mats[cur_mat][i];
        }
        out.push_back(std::move(m));
        cur_faces.clear();
    };

    std::istringstream in(text);
    std::string line;
    int lineno = 0;
    while (std::getline(in, line)) {
        lineno++;
        std::vector<std::string> tok = split_ws(line);
        if (tok.empty() || tok[0][0] == '#') continue;
        const std::string &k = tok[0];
        if (k == "v") {
            if (tok.size() < 4) {
                err = path + ":" + std::to_string(lineno) + ": vertex with fewer than 3 coordinates";
                return PATHS_ERR_PARSE;
            }
            for (int i = 1; i <= 3; i++) pos.push_back(strtof(tok[i].c_str(), nullptr));
        } else if (k == "f") {
            std::vector<int64_t> idx;
            for (size_t i = 1; i < tok.size(); i++) {
                long long vi = strtoll(tok[i].c_str(), nullptr, 10); // "v", "v/vt", "v//vn", "v/vt/vn": position index first
                int64_t n = (int64_t)(pos.size() / 3);
                int64_t ix = vi > 0 ? vi - 1 : n + vi;
                if (vi == 0 || ix < 0 || ix >= n) {
                    err = path + ":" + std::to_string(lineno) + ": face index out of range";
                    return PATHS_ERR_PARSE;
                }
                idx.push_back(ix);
            }
            for (size_t i = 1; i + 1 < idx.size(); i++) { // fan triangulation
                cur_faces.push_back(idx[0]);
                cur_faces.push_back(idx[i]);
                cur_faces.push_back(idx[i + 1]);
            }
        } else if (k == "o" || k == "g") {
            flush();
        } else if (k == "usemtl") {
            std::string name = join_from(tok, 1);
            if (!cur_faces.empty() && (!have_mat || name != cur_mat)) flush();
            cur_mat = name;
            have_mat = true;
        } else if (k == "mtllib") {
            std::string mtext;
            if (read_file(dir_of(path) + "/" + join_from(tok, 1), mtext)) { // a missing MTL is not fatal (src/obj.rs:12-21)
                std::istringstream min(mtext);
                std::string ml, cur;
                while (std::getline(min, ml)) {
                    std::vector<std::string> mt = split_ws(ml);
                    if (mt.empty() || mt[0][0] == '#') continue;
                    if (mt[0] == "newmtl") {
                        cur = join_from(mt, 1);
                        mats[cur] = {0.0, 0.0, 0.0};
                    } else if (mt[0] == "Kd" && !cur.empty() && mt.size() >= 4) {
                        mats[cur] = {(double)strtof(mt[1].c_str(), nullptr), (double)strtof(mt[2].c_str(), nullptr), (double)strtof(mt[3].c_str(), nullptr)};
                    }
                }
            }
        }
    }
    flush();
    return status;
}

// ------------------------------------------------------------------ PLY (src/ply.rs:11-78)
namespace {
int ply_size(const std::string &t) {
    if (t == "char" || t == "int8" || t == "uchar" || t == "uint8") return 1;
    if (t == "short" || t == "int16" || t == "ushort" || t == "uint16") return 2;
    if (t == "int" || t == "int32" || t == "uint" || t == "uint32" || t == "float" || t == "float32") return 4;
    if (t == "double" || t == "float64") return 8;
    return 0;
}
double ply_read(const unsigned char *p, const std::string &t, bool big) {
    unsigned char b[8];
    int n = ply_size(t);
    for (int i = 0; i < n; i++) b[i] = big ? p[n - 1 - i] : p[i];
    if (t == "char" || t == "int8") return (double)(int8_t)b[0];
    if (t == "uchar" || t == "uint8") return (double)b[0];
    if (t == "short" || t == "int16") { int16_t v; memcpy(&v, b, 2); return v; }
    if (t == "ushort" || t == "uint16") { uint16_t v; memcpy(&v, b, 2); return v; }
    if (t == "int" || t == "int32") { int32_t v; memcpy(&v, b, 4); return v; }
    if (t == "uint" || t == "uint32") { uint32_t v; memcpy(&v, b, 4); return v; }
    if (t == "float" || t == "float32") { float v; memcpy(&v, b, 4); return v; }
    double v;
    memcpy(&v, b, 8);
    return v;
}
struct PlyProp {
    bool list;
    std::string name, type, count_type;
};
struct PlyElem {
    std::string name;
    size_t count;
    std::vector<PlyProp> props;
};
} // namespace

paths_status load_ply(const std::string &path, LoadedModel &out, std::string &err) {
    std::string data;
    if (!read_file(path, data)) {
        err = "cannot open " + path;
        return PATHS_ERR_IO;
    }
    size_t eh = data.find("end_header");
    if (data.compare(0, 3, "ply") != 0 || eh == std::string::npos) {
        err = path + ": not a PLY file";
        return PATHS_ERR_PARSE;
    }
    size_t body = data.find('\n', eh);
    if (body == std::string::npos) {
        err = path + ": truncated header";
        return PATHS_ERR_PARSE;
    }
    body++;
    std::string fmt;
    std::vector<PlyElem> elems;
    {
        std::istringstream hs(data.substr(0, body));
        std::string line;
        while (std::getline(hs, line)) {
            std::vector<std::string> tok = split_ws(line);
            if (tok.empty()) continue;
            if (tok[0] == "format" && tok.size() >= 2) fmt = tok[1];
            else if (tok[0] == "element" && tok.size() >= 3) elems.push_back(PlyElem{tok[1], (size_t)strtoull(tok[2].c_str(), nullptr, 10), {}});
            else if (tok[0] == "property" && !elems.empty()) {
                if (tok.size() >= 5 && tok[1] == "list") elems.back().props.push_back(PlyProp{true, tok[4], tok[3], tok[2]});
                else if (tok.size() >= 3) elems.back().props.push_back(PlyProp{false, tok[2], tok[1], ""});
            }
        }
    }
    bool ascii = fmt == "ascii", big = fmt == "binary_big_endian";
    if (!ascii && !big && fmt != "binary_little_endian") {
        err = path + ": unknown PLY format '" + fmt + "'";
        return PATHS_ERR_PARSE;
    }
    out = LoadedModel();
    const unsigned char *p = (const unsigned char *)data.data() + body, *end = (const unsigned char *)data.data() + data.size();
    const char *ap = data.c_str() + body; // ascii cursor
    auto next_ascii = [&](double &v) -> bool {
        char *e = nullptr;
        v = strtod(ap, &e);
        if (e == ap) return false;
        ap = e;
        return true;
    };
    for (const PlyElem &el : elems) {
        bool is_vertex = el.name == "vertex", is_face = el.name == "face";
        int ix = -1, iy = -1, iz = -1, ir = -1, ig = -1, ib = -1;
        for (size_t k = 0; k < el.props.size(); k++) {
            const std::string &n = el.props[k].name;
            if (n == "x") ix = (int)k; else if (n == "y") iy = (int)k; else if (n == "z") iz = (int)k;
            else if (n == "red") ir = (int)k; else if (n == "green") ig = (int)k; else if (n == "blue") ib = (int)k;
        }
        if (is_vertex && (ix < 0 || iy < 0 || iz < 0)) {
            err = path + ": vertex element without x/y/z";
            return PATHS_ERR_PARSE;
        }
        bool colours = is_vertex && ir >= 0 && ig >= 0 && ib >= 0; // src/ply.rs:59-71
        // a header-declared count is checked against what the rest of the file can hold before anything loops over it or allocates:
        // every row needs at least one byte per property (ascii: a character; binary: the smallest scalar / list count)
        {
            const size_t left = ascii ? (size_t)((const char *)end - ap) : (size_t)(end - p);
            if (el.props.empty() ? el.count != 0 : el.count > left / el.props.size() + 1) {
                err = path + ": element '" + el.name + "' declares more rows than the file can hold";
                return PATHS_ERR_PARSE;
            }
        }
        std::vector<double> vals(el.props.size());
        std::vector<double> list;
        for (size_t row = 0; row < el.count; row++) {
            list.clear();
            for (size_t k = 0; k < el.props.size(); k++) {
                const PlyProp &pr = el.props[k];
                if (!pr.list) {
                    if (ascii) {
                        if (!next_ascii(vals[k])) { err = path + ": truncated ascii body"; return PATHS_ERR_PARSE; }
                    } else {
                        int n = ply_size(pr.type);
                        if (n == 0 || p + n > end) { err = path + ": truncated binary body"; return PATHS_ERR_PARSE; }
                        vals[k] = ply_read(p, pr.type, big);
                        p += n;
                    }
                } else {
                    double cnt;
                    if (ascii) {
                        if (!next_ascii(cnt)) { err = path + ": truncated ascii body"; return PATHS_ERR_PARSE; }
                    } else {
                        int n = ply_size(pr.count_type);
                        if (n == 0 || p + n > end) { err = path + ": truncated binary body"; return PATHS_ERR_PARSE; }
                        cnt = ply_read(p, pr.count_type, big);
                        p += n;
                    }
                    bool keep = is_face && (pr.name == "vertex_indices" || pr.name == "vertex_index");
                    if (!(cnt >= 0.0 && cnt <= 1048576.0)) { err = path + ": list length out of range"; return PATHS_ERR_PARSE; }
                    for (int c = 0; c < (int)cnt; c++) {
                        double v;
                        if (ascii) {
                            if (!next_ascii(v)) { err = path + ": truncated ascii body"; return PATHS_ERR_PARSE; }
                        } else {
                            int n = ply_size(pr.type);
                            if (n == 0 || p + n > end) { err = path + ": truncated binary body"; return PATHS_ERR_PARSE; }
                            v = ply_read(p, pr.type, big);
                            p += n;
                        }
                        if (keep) list.push_back(v);
                    }
                }
            }
            if (is_vertex) {
                out.vertices.push_back((double)(float)vals[ix]); // src/ply.rs:46-52 reads f32
                out.vertices.push_back((double)(float)vals[iy]);
                out.vertices.push_back((double)(float)vals[iz]);
                if (colours) {
                    out.colours.push_back(vals[ir] / 255.0);
                    out.colours.push_back(vals[ig] / 255.0);
                    out.colours.push_back(vals[ib] / 255.0);
                }
            } else if (is_face) {
                if (list.size() < 3) { err = path + ": face with fewer than 3 indices"; return PATHS_ERR_PARSE; }
                for (int c = 0; c < 3; c++) { // src/ply.rs:72-75 keeps the first three
                    if (!(list[c] >= 0.0 && list[c] <= 4294967294.0)) { err = path + ": face index out of range"; return PATHS_ERR_PARSE; }
                    out.faces.push_back((uint32_t)list[c]);
                }
            }
        }
    }
    return PATHS_OK;
}

// ------------------------------------------------------------------ Radiance HDR (image::hdr::HdrDecoder::read_image_hdr)
paths_status load_hdr(const std::string &path, uint32_t &w, uint32_t &h, std::vector<float> &texels, std::string &err) {
    std::string data;
    if (!read_file(path, data)) {
        err = "cannot open " + path;
        return PATHS_ERR_IO;
    }
    size_t pos = 0;
    w = h = 0;
    bool found = false;
    while (pos < data.size()) {
        size_t nl = data.find('\n', pos);
        if (nl == std::string::npos) break;
        std::vector<std::string> tok = split_ws(data.substr(pos, nl - pos));
        pos = nl + 1;
        if (tok.size() == 4 && (tok[0] == "-Y" || tok[0] == "+Y") && (tok[2] == "+X" || tok[2] == "-X")) {
            h = (uint32_t)strtoul(tok[1].c_str(), nullptr, 10);
            w = (uint32_t)strtoul(tok[3].c_str(), nullptr, 10);
            found = true;
            break;
        }
    }
    if (!found || w == 0 || h == 0) {
        err = path + ": no Radiance resolution line";
        return PATHS_ERR_PARSE;
    }
    // the resolution line is checked against the bytes that follow before anything is allocated: a flat scanline takes 4 bytes per
    // pixel, a run-length encoded one at least 4 + 4 * ceil(w / 127) * 2
    {
        const size_t left = data.size() - pos;
        const size_t min_row = std::min<size_t>((size_t)w * 4, 4 + 8 * (((size_t)w + 126) / 127));
        if ((size_t)h > left / std::max<size_t>(min_row, 1) + 1 || (uint64_t)w * h > (1ull << 31)) {
            err = path + ": resolution line declares more pixels than the file can hold";
            return PATHS_ERR_PARSE;
        }
    }
    std::vector<unsigned char> rgbe((size_t)w * h * 4);
    const unsigned char *d = (const unsigned char *)data.data();
    size_t n = data.size();
    for (uint32_t y = 0; y < h; y++) {
        unsigned char *row = &rgbe[(size_t)y * w * 4];
        if (w >= 8 && w <= 0x7FFF && pos + 4 <= n && d[pos] == 2 && d[pos + 1] == 2 && (d[pos + 2] & 0x80) == 0) {
            pos += 4; // new-style RLE: the four channels are stored one after another
            for (int c = 0; c < 4; c++) {
                uint32_t x = 0;
                while (x < w) {
                    if (pos >= n) { err = path + ": truncated RLE data"; return PATHS_ERR_PARSE; }
                    unsigned cnt = d[pos++];
                    if (cnt > 128) {
                        cnt -= 128;
                        if (pos >= n || x + cnt > w) { err = path + ": bad RLE run"; return PATHS_ERR_PARSE; }
                        unsigned char v = d[pos++];
                        for (unsigned k = 0; k < cnt; k++) row[4 * (x + k) + c] = v;
                    } else {
                        if (pos + cnt > n || x + cnt > w || cnt == 0) { err = path + ": bad RLE literal"; return PATHS_ERR_PARSE; }
                        for (unsigned k = 0; k < cnt; k++) row[4 * (x + k) + c] = d[pos + k];
                        pos += cnt;
                    }
                    x += cnt;
                }
            }
        } else {
            if (pos + (size_t)w * 4 > n) { err = path + ": truncated pixel data"; return PATHS_ERR_PARSE; }
            memcpy(row, d + pos, (size_t)w * 4);
            pos += (size_t)w * 4;
        }
    }
    texels.resize((size_t)w * h * 3);
    for (size_t i = 0; i < (size_t)w * h; i++) {
        int e = rgbe[4 * i + 3];
        float scale = e == 0 ? 0.0f : std::ldexp(1.0f, e - 136); // image 0.23: exp2(e - 128 - 8)
        texels[3 * i] = (float)rgbe[4 * i] * scale;
        texels[3 * i + 1] = (float)rgbe[4 * i + 1] * scale;
        texels[3 * i + 2] = (float)rgbe[4 * i + 2] * scale;
    }
    return PATHS_OK;
}

} // namespace paths
