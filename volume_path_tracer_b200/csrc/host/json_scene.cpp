// json_scene.cpp — scene/configuration JSON reader for the reference's schema.
//
// Replaces read_configuration (src/configuration.cpp:8-22), which uses glaze reflection over
// vpt::Configuration (include/vpt/configuration.hpp:14-65) with error_on_missing_keys = true: EVERY key
// of the schema is mandatory, unknown keys are an error, vectors are JSON arrays, output_size /
// tile_size / coord are [x, y]. glaze is not vendored with the reference, so this is our own reader:
// a small recursive-descent JSON parser + a schema walk.
#include "json_scene.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <vector>

namespace vpt_host {
namespace {

struct JValue {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<JValue> arr;
    std::vector<std::pair<std::string, JValue>> obj;
    const JValue* find(const std::string& k) const
    {
        for (const auto& kv : obj) if (kv.first == k) return &kv.second;
        return nullptr;
    }
};

struct Parser {
    const char* p; const char* end; std::string err;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
    bool fail(const std::string& m) { if (err.empty()) err = m + " at offset " + std::to_string(offset()); return false; }
    const char* begin0 = nullptr;
    size_t offset() const { return size_t(p - begin0); }
    bool parse_string(std::string& out)
    {
        if (p >= end || *p != '"') return fail("expected string");
        ++p; out.clear();
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) return fail("bad escape");
                switch (*p) {
                case '"': out += '"'; break; case '\\': out += '\\'; break; case '/': out += '/'; break;
                case 'b': out += '\b'; break; case 'f': out += '\f'; break; case 'n': out += '\n'; break;
                case 'r': out += '\r'; break; case 't': out += '\t'; break;
                case 'u': {
                    if (end - p < 5) return fail("bad \\u escape");
                    unsigned cp = unsigned(std::strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16));
                    p += 4;
                    if (cp < 0x80) out += char(cp);
                    else if (cp < 0x800) { out += char(0xC0 | (cp >> 6)); out += char(0x80 | (cp & 0x3F)); }
                    else { out += char(0xE0 | (cp >> 12)); out += char(0x80 | ((cp >> 6) & 0x3F)); out += char(0x80 | (cp & 0x3F)); }
                    break;
                }
                default: return fail("bad escape");
                }
                ++p;
            } else out += *p++;
        }
        if (p >= end) return fail("unterminated string");
        ++p;
        return true;
    }
    bool parse_value(JValue& v, int depth = 0)
    {
        if (depth > 64) return fail("nesting too deep");
        ws();
        if (p >= end) return fail("unexpected end of input");
        if (*p == '{') {
            v.kind = JValue::Object; ++p; ws();
            if (p < end && *p == '}') { ++p; return true; }
            for (;;) {
                ws();
                std::string k;
                if (!parse_string(k)) return false;
                ws();
                if (p >= end || *p != ':') return fail("expected ':'");
                ++p;
                JValue child;
                if (!parse_value(child, depth + 1)) return false;
                v.obj.emplace_back(std::move(k), std::move(child));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (*p == '[') {
            v.kind = JValue::Array; ++p; ws();
            if (p < end && *p == ']') { ++p; return true; }
            for (;;) {
                JValue child;
                if (!parse_value(child, depth + 1)) return false;
                v.arr.push_back(std::move(child));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (*p == '"') { v.kind = JValue::String; return parse_string(v.str); }
        if (end - p >= 4 && std::string(p, p + 4) == "true") { v.kind = JValue::Bool; v.b = true; p += 4; return true; }
        if (end - p >= 5 && std::string(p, p + 5) == "false") { v.kind = JValue::Bool; v.b = false; p += 5; return true; }
        if (end - p >= 4 && std::string(p, p + 4) == "null") { v.kind = JValue::Null; p += 4; return true; }
        char* e = nullptr;
        v.num = std::strtod(p, &e);
        if (e == p) return fail("unexpected character");
        v.kind = JValue::Number; p = e;
        return true;
    }
};

struct Schema {
    std::string err;
    bool fail(const std::string& m) { if (err.empty()) err = m; return false; }
    const JValue* key(const JValue& o, const std::string& path, const char* k)
    {
        const JValue* v = o.find(k);
        if (!v) fail("missing key \"" + path + k + "\"");
        return v;
    }
    bool only(const JValue& o, const std::string& path, std::initializer_list<const char*> allowed)
    {
        if (o.kind != JValue::Object) return fail("\"" + path + "\" must be an object");
        for (const auto& kv : o.obj) {
            bool ok = false;
            for (const char* a : allowed) ok |= kv.first == a;
            if (!ok) return fail("unknown key \"" + path + kv.first + "\"");
        }
        return true;
    }
    bool f32(const JValue& o, const std::string& path, const char* k, float& out)
    {
        const JValue* v = key(o, path, k);
        if (!v) return false;
        if (v->kind != JValue::Number) return fail("\"" + path + k + "\" must be a number");
        out = float(v->num);
        return true;
    }
    bool u32(const JValue& o, const std::string& path, const char* k, uint32_t& out)
    {
        const JValue* v = key(o, path, k);
        if (!v) return false;
        if (v->kind != JValue::Number || v->num < 0 || v->num > 4294967295.0 || v->num != std::floor(v->num))
            return fail("\"" + path + k + "\" must be an unsigned integer");
        out = uint32_t(v->num);
        return true;
    }
    bool boolean(const JValue& o, const std::string& path, const char* k, int32_t& out)
    {
        const JValue* v = key(o, path, k);
        if (!v) return false;
        if (v->kind != JValue::Bool) return fail("\"" + path + k + "\" must be true or false");
        out = v->b ? 1 : 0;
        return true;
    }
    bool vec3(const JValue& o, const std::string& path, const char* k, float out[3])
    {
        const JValue* v = key(o, path, k);
        if (!v) return false;
        if (v->kind != JValue::Array || v->arr.size() != 3) return fail("\"" + path + k + "\" must be an array of 3 numbers");
        for (int i = 0; i < 3; ++i) {
            if (v->arr[i].kind != JValue::Number) return fail("\"" + path + k + "\" must be an array of 3 numbers");
            out[i] = float(v->arr[i].num);
        }
        return true;
    }
    bool ivec2(const JValue& o, const std::string& path, const char* k, int64_t out[2])
    {
        const JValue* v = key(o, path, k);
        if (!v) return false;
        if (v->kind != JValue::Array || v->arr.size() != 2) return fail("\"" + path + k + "\" must be an array of 2 integers");
        for (int i = 0; i < 2; ++i) {
            // finite and exactly representable before the cast (int64_t(1e300) or int64_t(inf) is undefined behaviour)
            if (v->arr[i].kind != JValue::Number || !std::isfinite(v->arr[i].num) || std::fabs(v->arr[i].num) >= 9007199254740992.0 || v->arr[i].num != std::floor(v->arr[i].num))
                return fail("\"" + path + k + "\" must be an array of 2 integers");
            out[i] = int64_t(v->arr[i].num);
        }
        return true;
    }
};

} // namespace

bool read_text_file(const std::string& path, std::string& out, std::string& err)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open \"" + path + "\""; return false; }
    out.clear();
    char buf[65536];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) out.append(buf, n);
    std::fclose(f);
    return true;
}

bool parse_scene_json(const std::string& text, vptc_scene& s, std::string& volume_path, std::string& err)
{
    Parser ps{text.data(), text.data() + text.size(), {}};
    ps.begin0 = text.data();
    JValue root;
    if (!ps.parse_value(root)) { err = "JSON syntax error: " + ps.err; return false; }
    ps.ws();
    if (ps.p != ps.end) { err = "JSON syntax error: trailing characters at offset " + std::to_string(ps.offset()); return false; }

    Schema sc;
    bool ok = sc.only(root, "", {"seed", "output_size", "tile_size", "num_waves", "num_workers", "camera_parameters", "worker_parameters",
                                 "volume_path", "volume_parameters"});
    ok = ok && sc.u32(root, "", "seed", s.seed) && sc.ivec2(root, "", "output_size", s.output_size) && sc.ivec2(root, "", "tile_size", s.tile_size) &&
         sc.u32(root, "", "num_waves", s.num_waves) && sc.u32(root, "", "num_workers", s.num_workers);
    if (ok) {
        const JValue* v = sc.key(root, "", "volume_path");
        if (!v) ok = false;
        else if (v->kind != JValue::String) ok = sc.fail("\"volume_path\" must be a string");
        else volume_path = v->str;
    }
    if (ok) {
        const JValue* c = sc.key(root, "", "camera_parameters");
        ok = c && sc.only(*c, "camera_parameters.", {"position", "look", "up", "vfov_deg", "imaging_ratio"}) &&
             sc.vec3(*c, "camera_parameters.", "position", s.camera_parameters.position) && sc.vec3(*c, "camera_parameters.", "look", s.camera_parameters.look) &&
             sc.vec3(*c, "camera_parameters.", "up", s.camera_parameters.up) && sc.f32(*c, "camera_parameters.", "vfov_deg", s.camera_parameters.vfov_deg) &&
             sc.f32(*c, "camera_parameters.", "imaging_ratio", s.camera_parameters.imaging_ratio);
    }
    if (ok) {
        const JValue* w = sc.key(root, "", "worker_parameters");
        const std::string P = "worker_parameters.";
        ok = w && sc.only(*w, P, {"single_pixel", "use_jitter", "infinite_light", "distant_light", "max_depth"}) &&
             sc.boolean(*w, P, "use_jitter", s.worker_parameters.use_jitter) && sc.u32(*w, P, "max_depth", s.worker_parameters.max_depth);
        if (ok) {
            const JValue* sp = sc.key(*w, P, "single_pixel");
            ok = sp && sc.only(*sp, P + "single_pixel.", {"enabled", "coord"}) && sc.boolean(*sp, P + "single_pixel.", "enabled", s.worker_parameters.single_pixel_enabled) &&
                 sc.ivec2(*sp, P + "single_pixel.", "coord", s.worker_parameters.single_pixel_coord);
        }
        if (ok) {
            const JValue* il = sc.key(*w, P, "infinite_light");
            ok = il && sc.only(*il, P + "infinite_light.", {"xyz", "multiplier"}) && sc.vec3(*il, P + "infinite_light.", "xyz", s.worker_parameters.infinite_light.xyz) &&
                 sc.f32(*il, P + "infinite_light.", "multiplier", s.worker_parameters.infinite_light.multiplier);
        }
        if (ok) {
            const JValue* dl = sc.key(*w, P, "distant_light");
            ok = dl && sc.only(*dl, P + "distant_light.", {"xyz", "multiplier", "inv_direction"}) && sc.vec3(*dl, P + "distant_light.", "xyz", s.worker_parameters.distant_light.xyz) &&
                 sc.f32(*dl, P + "distant_light.", "multiplier", s.worker_parameters.distant_light.multiplier) &&
                 sc.vec3(*dl, P + "distant_light.", "inv_direction", s.worker_parameters.distant_light.inv_direction);
        }
    }
    if (ok) {
        const JValue* v = sc.key(root, "", "volume_parameters");
        const std::string P = "volume_parameters.";
        ok = v && sc.only(*v, P, {"henyey_greenstein_g", "le_scale", "sigma_a", "sigma_s", "temperature_offset", "temperature_scale"}) &&
             sc.f32(*v, P, "henyey_greenstein_g", s.volume_parameters.henyey_greenstein_g) && sc.f32(*v, P, "le_scale", s.volume_parameters.le_scale) &&
             sc.f32(*v, P, "sigma_a", s.volume_parameters.sigma_a) && sc.f32(*v, P, "sigma_s", s.volume_parameters.sigma_s) &&
             sc.f32(*v, P, "temperature_offset", s.volume_parameters.temperature_offset) && sc.f32(*v, P, "temperature_scale", s.volume_parameters.temperature_scale);
    }
    if (ok) {
        if (s.output_size[0] <= 0 || s.output_size[1] <= 0) ok = sc.fail("\"output_size\" must be positive");
        else if (s.tile_size[0] <= 0 || s.tile_size[1] <= 0) ok = sc.fail("\"tile_size\" must be positive");
    }
    if (!ok) { err = sc.err; return false; }
    return true;
}

} // namespace vpt_host
