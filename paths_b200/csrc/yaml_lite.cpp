// yaml_lite.cpp -- the YAML subset the reference's scene files use (what serde_yaml 0.8.11 is asked to read in
// src/main.rs:46): block mappings and sequences nested by indentation, flow mappings `{ x: 1, y: 2 }`, flow sequences
// `[]` / `[a, b]`, plain / single- / double-quoted scalars, `#` comments.  Anchors, tags, multi-documents and block
// scalars are not used by any scene and are rejected.
#include <cctype>
#include <cstdlib>

#include "loaders.h"

namespace paths {

const YamlNode *YamlNode::get(const std::string &key) const {
    if (kind != Map) return nullptr;
    for (const auto &kv : map)
        if (kv.first == key) return kv.second.get();
    return nullptr;
}

namespace {

struct Line {
    int indent;
    std::string text; // comment-stripped, right-trimmed, without the indentation
    int number;
};

std::string rtrim(const std::string &s) {
    size_t e = s.size();
    while (e > 0 && (s[e - 1] == ' ' || s[e - 1] == '\t' || s[e - 1] == '\r')) e--;
    return s.substr(0, e);
}
std::string trim(const std::string &s) {
    size_t b = 0;
    while (b < s.size() && (s[b] == ' ' || s[b] == '\t')) b++;
    return rtrim(s.substr(b));
}

// removes a trailing comment: '#' at line start or preceded by whitespace, outside quotes
std::string strip_comment(const std::string &s) {
    char q = 0;
    for (size_t i = 0; i < s.size(); i++) {
        char c = s[i];
        if (q) {
            if (c == q) q = 0;
            else if (c == '\\' && q == '"') i++;
        } else if (c == '"' || c == '\'') q = c;
        else if (c == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
    }
    return s;
}

struct Parser {
    std::vector<Line> lines;
    std::string err;

    bool fail(int line, const std::string &msg) {
        err = "yaml line " + std::to_string(line) + ": " + msg;
        return false;
    }

    // position of the ':' that ends a mapping key in `s`, or npos
    static size_t key_colon(const std::string &s) {
        char q = 0;
        int depth = 0;
        for (size_t i = 0; i < s.size(); i++) {
            char c = s[i];
            if (q) {
                if (c == q) q = 0;
                else if (c == '\\' && q == '"') i++;
                continue;
            }
            if (c == '"' || c == '\'') q = c;
            else if (c == '{' || c == '[') depth++;
            else if (c == '}' || c == ']') depth--;
            else if (c == ':' && depth == 0 && (i + 1 == s.size() || s[i + 1] == ' ' || s[i + 1] == '\t')) return i;
        }
        return std::string::npos;
    }

    static std::string unquote(const std::string &s, bool &quoted) {
        quoted = false;
        if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\''))) {
            quoted = true;
            std::string body = s.substr(1, s.size() - 2), out;
            if (s.front() == '\'') {
                for (size_t i = 0; i < body.size(); i++) {
                    out += body[i];
                    if (body[i] == '\'' && i + 1 < body.size() && body[i + 1] == '\'') i++;
                }
            } else {
                for (size_t i = 0; i < body.size(); i++) {
                    if (body[i] == '\\' && i + 1 < body.size()) {
                        char n = body[++i];
                        out += n == 'n' ? '\n' : n == 't' ? '\t' : n;
                    } else out += body[i];
                }
            }
            return out;
        }
        return s;
    }

    // flow / scalar value starting at s[pos]; advances pos past it
    bool parse_flow(const std::string &s, size_t &pos, YamlNode &out, int line) {
        while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\t')) pos++;
        out.line = line;
        if (pos >= s.size()) {
            out.kind = YamlNode::Null;
            return true;
        }
        if (s[pos] == '{') {
            out.kind = YamlNode::Map;
            pos++;
            for (;;) {
                while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\t' || s[pos] == ',')) pos++;
                if (pos >= s.size()) return fail(line, "unterminated flow mapping");
                if (s[pos] == '}') {
                    pos++;
                    return true;
                }
                size_t kstart = pos;
                char q = 0;
                while (pos < s.size()) {
                    char c = s[pos];
                    if (q) {
                        if (c == q) q = 0;
                    } else if (c == '"' || c == '\'') q = c;
                    else if (c == ':' || c == ',' || c == '}') break;
                    pos++;
                }
                if (pos >= s.size() || s[pos] != ':') return fail(line, "flow mapping entry without ':'");
                bool qd;
                std::string key = unquote(trim(s.substr(kstart, pos - kstart)), qd);
                pos++;
                auto child = std::make_shared<YamlNode>();
                if (!parse_flow(s, pos, *child, line)) return false;
                out.map.emplace_back(key, child);
            }
        }
        if (s[pos] == '[') {
            out.kind = YamlNode::Seq;
            pos++;
            for (;;) {
                while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\t' || s[pos] == ',')) pos++;
                if (pos >= s.size()) return fail(line, "unterminated flow sequence");
                if (s[pos] == ']') {
                    pos++;
                    return true;
                }
                auto child = std::make_shared<YamlNode>();
                if (!parse_flow(s, pos, *child, line)) return false;
                out.seq.push_back(child);
            }
        }
        // scalar: up to an unquoted ',', '}' or ']' (inside flow context) or the end
        size_t start = pos;
        char q = 0;
        while (pos < s.size()) {
            char c = s[pos];
            if (q) {
                if (c == q) q = 0;
                else if (c == '\\' && q == '"') pos++;
            } else if (c == '"' || c == '\'') q = c;
            else if (c == ',' || c == '}' || c == ']') break;
            pos++;
        }
        std::string tok = trim(s.substr(start, pos - start));
        if (!tok.empty() && (tok[0] == '&' || tok[0] == '*' || tok[0] == '!' || tok[0] == '|' || tok[0] == '>'))
            return fail(line, "anchors, tags and block scalars are not supported");
        out.scalar = unquote(tok, out.quoted);
        out.kind = (!out.quoted && (out.scalar.empty() || out.scalar == "~" || out.scalar == "null")) ? YamlNode::Null : YamlNode::Scalar;
        return true;
    }

    bool parse_inline(const std::string &s, YamlNode &out, int line) {
        size_t pos = 0;
        std::string t = trim(s);
        // a plain scalar outside flow context may contain ',' : only treat as flow when it starts with { or [
        if (!t.empty() && t[0] != '{' && t[0] != '[') {
            if (t[0] == '&' || t[0] == '*' || t[0] == '!' || t[0] == '|' || t[0] == '>') return fail(line, "anchors, tags and block scalars are not supported");
            out.line = line;
            out.scalar = unquote(t, out.quoted);
            out.kind = (!out.quoted && (out.scalar == "~" || out.scalar == "null")) ? YamlNode::Null : YamlNode::Scalar;
            return true;
        }
        if (!parse_flow(t, pos, out, line)) return false;
        while (pos < t.size() && (t[pos] == ' ' || t[pos] == '\t')) pos++;
        if (pos != t.size()) return fail(line, "trailing characters after flow collection");
        return true;
    }

    // parses the block that starts at lines[i] with indentation `indent`; leaves i at the first line not consumed
    bool parse_block(size_t &i, int indent, YamlNode &out) {
        if (i >= lines.size()) {
            out.kind = YamlNode::Null;
            return true;
        }
        out.line = lines[i].number;
        bool is_seq = lines[i].text == "-" || lines[i].text.compare(0, 2, "- ") == 0;
        if (is_seq) {
            out.kind = YamlNode::Seq;
            while (i < lines.size() && lines[i].indent == indent && (lines[i].text == "-" || lines[i].text.compare(0, 2, "- ") == 0)) {
                std::string rest = lines[i].text.size() > 1 ? lines[i].text.substr(1) : "";
                size_t off = 0;
                while (off < rest.size() && rest[off] == ' ') off++;
                rest = rest.substr(off);
                auto child = std::make_shared<YamlNode>();
                if (rest.empty()) {
                    i++;
                    if (i < lines.size() && lines[i].indent > indent) {
                        if (!parse_block(i, lines[i].indent, *child)) return false;
                    }
                } else if (rest[0] != '{' && rest[0] != '[' && rest[0] != '"' && rest[0] != '\'' && key_colon(rest) != std::string::npos) {
                    // "- key: value": a mapping whose first entry shares the dash's line
                    lines[i].indent = indent + 1 + (int)off;
                    lines[i].text = rest;
                    if (!parse_block(i, lines[i].indent, *child)) return false;
                } else {
                    if (!parse_inline(rest, *child, lines[i].number)) return false;
                    i++;
                }
                out.seq.push_back(child);
            }
            if (i < lines.size() && lines[i].indent > indent) return fail(lines[i].number, "bad indentation in sequence");
            return true;
        }
        out.kind = YamlNode::Map;
        while (i < lines.size() && lines[i].indent == indent) {
            const std::string &t = lines[i].text;
            if (t == "-" || t.compare(0, 2, "- ") == 0) break;
            size_t c = key_colon(t);
            if (c == std::string::npos) return fail(lines[i].number, "expected 'key: value'");
            bool qd;
            std::string key = unquote(trim(t.substr(0, c)), qd);
            std::string rest = trim(t.substr(c + 1));
            auto child = std::make_shared<YamlNode>();
            int ln = lines[i].number;
            if (rest.empty()) {
                i++;
                if (i < lines.size() && (lines[i].indent > indent ||
                                         (lines[i].indent == indent && (lines[i].text == "-" || lines[i].text.compare(0, 2, "- ") == 0)))) {
                    if (!parse_block(i, lines[i].indent, *child)) return false;
                } else {
                    child->kind = YamlNode::Null;
                    child->line = ln;
                }
            } else {
                if (!parse_inline(rest, *child, ln)) return false;
                i++;
            }
            for (const auto &kv : out.map)
                if (kv.first == key) return fail(ln, "duplicate key '" + key + "'");
            out.map.emplace_back(key, child);
        }
        if (i < lines.size() && lines[i].indent > indent) return fail(lines[i].number, "bad indentation in mapping");
        return true;
    }
};

} // namespace

bool parse_yaml(const std::string &text, YamlNode &root, std::string &err) {
    Parser p;
    size_t pos = 0;
    int number = 0;
    while (pos <= text.size()) {
        size_t nl = text.find('\n', pos);
        if (nl == std::string::npos) nl = text.size();
        std::string raw = text.substr(pos, nl - pos);
        pos = nl + 1;
        number++;
        std::string s = rtrim(strip_comment(raw));
        size_t ind = 0;
        while (ind < s.size() && s[ind] == ' ') ind++;
        if (ind < s.size() && s[ind] == '\t') {
            err = "yaml line " + std::to_string(number) + ": tab used for indentation";
            return false;
        }
        if (ind == s.size()) continue;
        if (s.compare(ind, 3, "---") == 0 && p.lines.empty()) continue; // a leading document marker
        p.lines.push_back(Line{(int)ind, s.substr(ind), number});
    }
    root = YamlNode();
    if (p.lines.empty()) return true;
    size_t i = 0;
    if (!p.parse_block(i, p.lines[0].indent, root)) {
        err = p.err;
        return false;
    }
    if (i != p.lines.size()) {
        err = "yaml line " + std::to_string(p.lines[i].number) + ": unexpected de-indentation";
        return false;
    }
    return true;
}

} // namespace paths
