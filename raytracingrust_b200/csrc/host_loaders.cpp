// host_loaders.cpp — Scene::from_file (lib/core/scene.rs:74-76) and Mesh::from_file
// (lib/core/mesh.rs:63-76) for the C ABI.  Pure host code, no CUDA.
//
// JSON: the serde_json/typetag format the crate writes ("type" tag on every dyn Object /
// Material / Background, traits.rs:17,22,27).  Numbers are parsed as f64 and narrowed to f32,
// as serde_json does.  `settings.clamp_indirect` is absent from the shipped scenes although the
// struct has no serde default (scene.rs:23-30); it defaults to 10.0 (Scene::default, :55).
// OBJ: obj-rs `load_obj::<obj::Vertex, u32>` semantics: triangles with normals only,
// (position, normal) index pairs de-duplicated in first-use order.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "../../include/rtr.h"

namespace {

// ------------------------------------------------------------------ minimal JSON
struct JValue {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<JValue> arr;
    std::vector<std::pair<std::string, JValue>> obj;
    const JValue* get(const char* key) const {
        for (const auto& kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
};

struct JParser {
    const char* p;
    const char* end;
    std::string err;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
    bool fail(const char* m) { if (err.empty()) err = m; return false; }
    bool parse_string(std::string& out) {
        if (p >= end || *p != '"') return fail("expected string");
        ++p;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) return fail("bad escape");
                switch (*p) {
                    case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break; case 'f': out += '\f'; break;
                    case 'u': { if (end - p < 5) return fail("bad \\u"); out += '?'; p += 4; break; }
                    default: out += *p;
                }
                ++p;
            } else out += *p++;
        }
        if (p >= end) return fail("unterminated string");
        ++p;
        return true;
    }
    bool parse(JValue& v, int depth = 0) {
        if (depth > 256) return fail("nesting too deep");
        ws();
        if (p >= end) return fail("unexpected end");
        if (*p == '{') {
            v.kind = JValue::Obj; ++p; ws();
            if (p < end && *p == '}') { ++p; return true; }
            for (;;) {
                ws();
                std::string key;
                if (!parse_string(key)) return false;
                ws();
                if (p >= end || *p != ':') return fail("expected ':'");
                ++p;
                JValue child;
                if (!parse(child, depth + 1)) return false;
                v.obj.emplace_back(std::move(key), std::move(child));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (*p == '[') {
            v.kind = JValue::Arr; ++p; ws();
            if (p < end && *p == ']') { ++p; return true; }
            for (;;) {
                JValue child;
                if (!parse(child, depth + 1)) return false;
                v.arr.push_back(std::move(child));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (*p == '"') { v.kind = JValue::Str; return parse_string(v.str); }
        if (end - p >= 4 && !std::strncmp(p, "true", 4)) { v.kind = JValue::Bool; v.b = true; p += 4; return true; }
        if (end - p >= 5 && !std::strncmp(p, "false", 5)) { v.kind = JValue::Bool; v.b = false; p += 5; return true; }
        if (end - p >= 4 && !std::strncmp(p, "null", 4)) { v.kind = JValue::Null; p += 4; return true; }
        char* e = nullptr;
        v.num = std::strtod(p, &e);
        if (e == p) return fail("unexpected token");
        v.kind = JValue::Num;
        p = e;
        return true;
    }
};

struct LoadError { std::string msg; };
[[noreturn]] void bail(const std::string& m) { throw LoadError{m}; }

const JValue& field(const JValue& o, const char* key) {
    if (o.kind != JValue::Obj) bail(std::string("expected an object holding '") + key + "'");
    const JValue* v = o.get(key);
    if (!v) bail(std::string("missing field `") + key + "`");  // serde's wording
    return *v;
}
float f32_of(const JValue& v) { if (v.kind != JValue::Num) bail("expected a number"); return (float)v.num; }
uint32_t u32_of(const JValue& v) {
    if (v.kind != JValue::Num || v.num < 0 || v.num != std::floor(v.num) || v.num > 4294967295.0) bail("expected an unsigned integer");
    return (uint32_t)v.num;
}
bool bool_of(const JValue& v) { if (v.kind != JValue::Bool) bail("expected a bool"); return v.b; }
rtr_vec3 vec3_of(const JValue& v) { return rtr_vec3{f32_of(field(v, "x")), f32_of(field(v, "y")), f32_of(field(v, "z"))}; }
rtr_vec3 color_of(const JValue& v) { return rtr_vec3{f32_of(field(v, "r")), f32_of(field(v, "g")), f32_of(field(v, "b"))}; }
const std::string& type_of(const JValue& v) {
    const JValue& t = field(v, "type");
    if (t.kind != JValue::Str) bail("\"type\" must be a string");
    return t.str;
}

}  // namespace

struct rtr_host_scene {
    rtr_scene_desc desc;
    rtr_camera camera;
    rtr_settings settings;
    std::vector<rtr_object_ref> objects;
    std::vector<rtr_sphere> spheres;
    std::vector<rtr_volume> volumes;
    std::vector<rtr_material> materials;
    std::vector<float> texels;
};

struct rtr_host_mesh {
    rtr_mesh mesh;
    std::vector<float> positions, normals;
    std::vector<uint32_t> indices;
};

namespace {

uint32_t parse_material(rtr_host_scene& hs, const JValue& v, int slot) {
    uint32_t idx;
    if (slot >= 0) idx = (uint32_t)slot;
    else { idx = (uint32_t)hs.materials.size(); hs.materials.push_back(rtr_material{}); }
    rtr_material m;
    std::memset(&m, 0, sizeof(m));
    const std::string& t = type_of(v);
    if (t == "Lambertian") { m.kind = RTR_MAT_LAMBERTIAN; m.color = color_of(field(v, "albedo")); }
    else if (t == "Metal") { m.kind = RTR_MAT_METAL; m.color = color_of(field(v, "albedo")); m.param = f32_of(field(v, "fuzz")); }
    else if (t == "Emission") { m.kind = RTR_MAT_EMISSION; m.color = color_of(field(v, "color")); }
    else if (t == "Dielectric") { m.kind = RTR_MAT_DIELECTRIC; m.param = f32_of(field(v, "ir")); }
    else if (t == "Isotropic") { m.kind = RTR_MAT_ISOTROPIC; m.color = color_of(field(v, "color")); }
    else if (t == "MixMaterial") {
        m.kind = RTR_MAT_MIX;
        m.param = f32_of(field(v, "factor"));
        m.first = parse_material(hs, field(v, "first"), -1);
        m.second = parse_material(hs, field(v, "second"), -1);
    } else bail("unknown variant `" + t + "` for Material");
    hs.materials[idx] = m;
    return idx;
}

rtr_object_ref parse_object(rtr_host_scene& hs, const JValue& v, bool allow_volume) {
    const std::string& t = type_of(v);
    if (t == "Sphere") {
        rtr_sphere s;
        s.center = vec3_of(field(v, "center"));
        s.radius = f32_of(field(v, "radius"));
        s.material = u32_of(field(v, "material"));
        hs.spheres.push_back(s);
        return rtr_object_ref{RTR_OBJ_SPHERE, (uint32_t)hs.spheres.size() - 1};
    }
    if (t == "Volume") {
        if (!allow_volume) throw LoadError{"unsupported: Volume bounded by a Volume"};
        rtr_volume vol;
        vol.boundary = parse_object(hs, field(v, "boundary"), false);
        vol.neg_inv_density = f32_of(field(v, "neg_inv_density"));
        hs.volumes.push_back(vol);
        return rtr_object_ref{RTR_OBJ_VOLUME, (uint32_t)hs.volumes.size() - 1};
    }
    if (t == "Mesh") bail("Mesh cannot be deserialised (the reference panics, mesh.rs:51-60)");
    bail("unknown variant `" + t + "` for Object");
}

}  // namespace

namespace rtr { void set_global_error(const std::string& m); }  // capi.cpp

extern "C" {

rtr_status rtr_scene_load_json(const char* path, rtr_host_scene** out) {
    if (!path || !out) { rtr::set_global_error("rtr_scene_load_json: null argument"); return RTR_INVALID_ARG; }
    *out = nullptr;
    FILE* f = std::fopen(path, "rb");
    if (!f) { rtr::set_global_error(std::string("cannot open ") + path); return RTR_IO_ERROR; }
    std::string text;
    char buf[1 << 16];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, n);
    std::fclose(f);
    JParser jp{text.data(), text.data() + text.size(), {}};
    JValue root;
    if (!jp.parse(root)) { rtr::set_global_error(std::string(path) + ": JSON error: " + jp.err); return RTR_INVALID_ARG; }
    std::unique_ptr<rtr_host_scene> hs(new rtr_host_scene());
    try {
        const JValue& cam = field(root, "camera");
        hs->camera.lookfrom = vec3_of(field(cam, "lookfrom"));
        hs->camera.lookat = vec3_of(field(cam, "lookat"));
        hs->camera.vertical = vec3_of(field(cam, "vertical"));
        hs->camera.vertical_fov = f32_of(field(cam, "vertical_fov"));
        hs->camera.aspect_ratio = f32_of(field(cam, "aspect_ratio"));
        const JValue& st = field(root, "settings");
        hs->settings.samples_per_pixel = u32_of(field(st, "samples_per_pixel"));
        hs->settings.max_ray_depth = u32_of(field(st, "max_ray_depth"));
        if (hs->settings.max_ray_depth > 255) bail("max_ray_depth does not fit u8");
        const JValue* clamp = st.get("clamp_indirect");
        hs->settings.clamp_indirect = clamp ? f32_of(*clamp) : 10.0f;
        hs->settings.enable_multithreading = bool_of(field(st, "enable_multithreading"));
        hs->settings.enable_bvh_tree = bool_of(field(st, "enable_bvh_tree"));
        const JValue& mode = field(st, "mode");
        if (mode.kind != JValue::Str) bail("mode must be a string");
        if (mode.str == "Full") hs->settings.mode = RTR_MODE_FULL;
        else if (mode.str == "Clay") hs->settings.mode = RTR_MODE_CLAY;
        else if (mode.str == "Random") hs->settings.mode = RTR_MODE_RANDOM;
        else if (mode.str == "Normal") hs->settings.mode = RTR_MODE_NORMAL;
        else bail("unknown variant `" + mode.str + "` for RenderMode");

        const JValue& mats = field(root, "materials");
        if (mats.kind != JValue::Arr) bail("materials must be an array");
        hs->materials.resize(mats.arr.size());  // handles keep their indices; mix children appended
        for (size_t i = 0; i < mats.arr.size(); ++i) parse_material(*hs, mats.arr[i], (int)i);
        const uint32_t n_handles = (uint32_t)mats.arr.size();

        const JValue& objs = field(root, "objects");
        if (objs.kind != JValue::Arr) bail("objects must be an array");
        for (const JValue& o : objs.arr) hs->objects.push_back(parse_object(*hs, o, true));
        for (const rtr_sphere& s : hs->spheres)
            if (s.material >= n_handles) bail("material handle out of range (index panic in the reference)");

        const JValue& bg = field(root, "background");
        std::memset(&hs->desc, 0, sizeof(hs->desc));
        rtr_background& b = hs->desc.background;
        const std::string& bt = type_of(bg);
        if (bt == "UniformBackground") { b.kind = RTR_BG_UNIFORM; b.color = color_of(field(bg, "color")); }
        else if (bt == "GradientBackground") { b.kind = RTR_BG_GRADIENT; b.color = color_of(field(bg, "top")); b.bottom = color_of(field(bg, "bottom")); }
        else if (bt == "SkyMap") {
            b.kind = RTR_BG_SKYMAP;
            b.width = u32_of(field(bg, "width")); b.height = u32_of(field(bg, "height"));
            const JValue& img = field(bg, "image");
            if (img.kind != JValue::Arr || img.arr.size() != (size_t)b.width * b.height || img.arr.empty()) bail("SkyMap image size mismatch");
            hs->texels.reserve(img.arr.size() * 3);
            for (const JValue& c : img.arr) { rtr_vec3 k = color_of(c); hs->texels.push_back(k.x); hs->texels.push_back(k.y); hs->texels.push_back(k.z); }
            b.texels = hs->texels.data();
        } else bail("unknown variant `" + bt + "` for Background");
    } catch (const LoadError& e) {
        rtr::set_global_error(std::string(path) + ": " + e.msg);
        return e.msg.rfind("unsupported", 0) == 0 ? RTR_UNSUPPORTED : RTR_INVALID_ARG;
    }
    rtr_scene_desc& d = hs->desc;
    d.objects = hs->objects.data(); d.n_objects = (uint32_t)hs->objects.size();
    d.spheres = hs->spheres.data(); d.n_spheres = (uint32_t)hs->spheres.size();
    d.meshes = nullptr; d.n_meshes = 0;
    d.volumes = hs->volumes.data(); d.n_volumes = (uint32_t)hs->volumes.size();
    d.materials = hs->materials.data(); d.n_materials = (uint32_t)hs->materials.size();
    *out = hs.release();
    return RTR_OK;
}

const rtr_scene_desc* rtr_host_scene_desc(const rtr_host_scene* hs) { return hs ? &hs->desc : nullptr; }
const rtr_camera* rtr_host_scene_camera(const rtr_host_scene* hs) { return hs ? &hs->camera : nullptr; }
const rtr_settings* rtr_host_scene_settings(const rtr_host_scene* hs) { return hs ? &hs->settings : nullptr; }
void rtr_host_scene_destroy(rtr_host_scene* hs) { delete hs; }

}  // extern "C"

// ------------------------------------------------------------------ OBJ (mesh.rs:63-76)
namespace {
// one "a/b/c" face corner -> (position index, normal index), 0-based; negative = relative
bool parse_corner(const char* s, const char* e, long np, long nn, long* pi, long* ni) {
    long vals[3] = {0, 0, 0};
    bool have[3] = {false, false, false};
    int k = 0;
    const char* q = s;
    while (q < e && k < 3) {
        const char* slash = q;
        while (slash < e && *slash != '/') ++slash;
        if (slash > q) { vals[k] = std::strtol(std::string(q, slash).c_str(), nullptr, 10); have[k] = true; }
        ++k;
        q = slash + 1;
    }
    if (!have[0] || !have[2]) return false;  // normals are required by obj::Vertex
    *pi = vals[0] > 0 ? vals[0] - 1 : np + vals[0];
    *ni = vals[2] > 0 ? vals[2] - 1 : nn + vals[2];
    return *pi >= 0 && *pi < np && *ni >= 0 && *ni < nn;
}
}  // namespace

extern "C" {

rtr_status rtr_mesh_load_obj(const char* path, uint32_t material, rtr_host_mesh** out) {
    if (!path || !out) { rtr::set_global_error("rtr_mesh_load_obj: null argument"); return RTR_INVALID_ARG; }
    *out = nullptr;
    FILE* f = std::fopen(path, "rb");
    if (!f) { rtr::set_global_error(std::string("cannot open ") + path); return RTR_IO_ERROR; }
    std::vector<float> pos, nrm;
    std::unique_ptr<rtr_host_mesh> hm(new rtr_host_mesh());
    std::map<std::pair<long, long>, uint32_t> cache;
    char line[4096];
    size_t lineno = 0;
    auto err = [&](const std::string& m) {
        std::fclose(f);
        rtr::set_global_error(std::string(path) + ":" + std::to_string(lineno) + ": " + m);
        return RTR_INVALID_ARG;
    };
    while (std::fgets(line, sizeof(line), f)) {
        ++lineno;
        char* s = line;
        while (*s == ' ' || *s == '\t') ++s;
        if (s[0] == 'v' && (s[1] == ' ' || s[1] == '\t')) {
            char* e = s + 1;
            for (int k = 0; k < 3; ++k) { char* n = nullptr; pos.push_back((float)std::strtod(e, &n)); if (n == e) return err("bad vertex"); e = n; }
        } else if (s[0] == 'v' && s[1] == 'n' && (s[2] == ' ' || s[2] == '\t')) {
            char* e = s + 2;
            for (int k = 0; k < 3; ++k) { char* n = nullptr; nrm.push_back((float)std::strtod(e, &n)); if (n == e) return err("bad normal"); e = n; }
        } else if (s[0] == 'f' && (s[1] == ' ' || s[1] == '\t')) {
            const long np = (long)(pos.size() / 3), nn = (long)(nrm.size() / 3);
            uint32_t corner_idx[3];
            int corners = 0;
            char* q = s + 1;
            for (;;) {
                while (*q == ' ' || *q == '\t') ++q;
                if (*q == '\0' || *q == '\n' || *q == '\r') break;
                char* e = q;
                while (*e && *e != ' ' && *e != '\t' && *e != '\n' && *e != '\r') ++e;
                if (corners == 3) return err("Model should be triangulated first to be loaded properly");
                long pi, ni;
                if (!parse_corner(q, e, np, nn, &pi, &ni)) return err("Tried to extract normal data which are not contained in the model");
                auto key = std::make_pair(pi, ni);
                auto it = cache.find(key);
                if (it == cache.end()) {
                    uint32_t idx = (uint32_t)(hm->positions.size() / 3);
                    for (int k = 0; k < 3; ++k) { hm->positions.push_back(pos[3 * pi + k]); hm->normals.push_back(nrm[3 * ni + k]); }
                    it = cache.emplace(key, idx).first;
                }
                corner_idx[corners++] = it->second;
                q = e;
            }
            if (corners != 3) return err("Model should be triangulated first to be loaded properly");
            for (int k = 0; k < 3; ++k) hm->indices.push_back(corner_idx[k]);
        }
    }
    std::fclose(f);
    hm->mesh.positions = hm->positions.data();
    hm->mesh.normals = hm->normals.data();
    hm->mesh.indices = hm->indices.data();
    hm->mesh.n_vertices = (uint32_t)(hm->positions.size() / 3);
    hm->mesh.n_indices = (uint32_t)hm->indices.size();
    hm->mesh.material = material;
    hm->mesh._pad = 0;
    *out = hm.release();
    return RTR_OK;
}

const rtr_mesh* rtr_host_mesh_get(const rtr_host_mesh* hm) { return hm ? &hm->mesh : nullptr; }
void rtr_host_mesh_destroy(rtr_host_mesh* hm) { delete hm; }

}  // extern "C"

// ------------------------------------------------------------------ headless output
// The reference only ever shows RenderTarget.data in a window (src/main.rs:92-97); without a graphics stack the
// RGBA8 buffer is written to a file instead.  Minimal PNG: 8-bit RGBA, filter 0 on every row, zlib deflate.
#include <zlib.h>

extern "C" rtr_status rtr_write_png(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height) {
    if (!path || (!rgba8 && width && height)) { rtr::set_global_error("rtr_write_png: null argument"); return RTR_INVALID_ARG; }
    if (width == 0 || height == 0) { rtr::set_global_error("rtr_write_png: empty image"); return RTR_INVALID_ARG; }
    std::vector<unsigned char> raw((size_t)height * ((size_t)width * 4 + 1));
    for (uint32_t y = 0; y < height; ++y) {
        unsigned char* row = &raw[(size_t)y * ((size_t)width * 4 + 1)];
        row[0] = 0;  // filter type None
        std::memcpy(row + 1, rgba8 + (size_t)y * width * 4, (size_t)width * 4);
    }
    uLongf zlen = compressBound((uLong)raw.size());
    std::vector<unsigned char> z(zlen);
    if (compress2(z.data(), &zlen, raw.data(), (uLong)raw.size(), 6) != Z_OK) { rtr::set_global_error("rtr_write_png: deflate failed"); return RTR_IO_ERROR; }
    FILE* f = std::fopen(path, "wb");
    if (!f) { rtr::set_global_error(std::string("rtr_write_png: cannot open ") + path); return RTR_IO_ERROR; }
    auto be32 = [](unsigned char* p, uint32_t v) { p[0] = (unsigned char)(v >> 24); p[1] = (unsigned char)(v >> 16); p[2] = (unsigned char)(v >> 8); p[3] = (unsigned char)v; };
    bool ok = true;
    auto chunk = [&](const char* type, const unsigned char* data, uint32_t len) {
        unsigned char hdr[8];
        be32(hdr, len); std::memcpy(hdr + 4, type, 4);
        uLong c = crc32(0L, (const Bytef*)hdr + 4, 4);
        if (len) c = crc32(c, data, len);
        unsigned char tail[4]; be32(tail, (uint32_t)c);
        ok = ok && std::fwrite(hdr, 1, 8, f) == 8 && (len == 0 || std::fwrite(data, 1, len, f) == len) && std::fwrite(tail, 1, 4, f) == 4;
    };
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    ok = std::fwrite(sig, 1, 8, f) == 8;
    unsigned char ihdr[13];
    be32(ihdr, width); be32(ihdr + 4, height);
    ihdr[8] = 8; ihdr[9] = 6; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;  // 8-bit, RGBA, deflate, adaptive, no interlace
    chunk("IHDR", ihdr, 13);
    chunk("IDAT", z.data(), (uint32_t)zlen);
    chunk("IEND", nullptr, 0);
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) { rtr::set_global_error("rtr_write_png: write failed"); return RTR_IO_ERROR; }
    return RTR_OK;
}
