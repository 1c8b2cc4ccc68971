// GLB (binary glTF 2.0) import with the reference's LoadGltf semantics (src/win32_main.cpp:1454-1593):
//   * depth-first walk over the scene nodes through a LIFO stack, node transforms IGNORED;
//   * only triangle primitives; POSITION + indices (or sequential vertices when un-indexed);
//   * one new material per primitive from pbrMetallicRoughness.baseColorFactor when the material has
//     no base-colour texture, otherwise material index 1 (main.cpp:1504-1515);
//   * glTF roughness / metallic / ior are ignored (material_t defaults: roughness 1, metalness 0).
// Container + JSON parsing are written from the glTF 2.0 / RFC 8259 specifications (the reference
// uses the third-party cgltf; none of it is used here).
#include "host_internal.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>

namespace pth {
namespace {

struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    const Json *get(const char *key) const {
        if (kind != Obj) return nullptr;
        for (auto &kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    long as_int(long dflt) const { return kind == Num ? (long)num : dflt; }
    size_t size() const { return kind == Arr ? arr.size() : 0; }
};

struct Parser {
    const char *p, *end;
    bool ok = true;
    int depth = 0;                    // nesting of arrays / objects; a hostile file must not overflow the C++ stack
    static constexpr int kMaxDepth = 64;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    bool lit(const char *s) { size_t n = std::strlen(s); if ((size_t)(end - p) >= n && !std::memcmp(p, s, n)) { p += n; return true; } return false; }
    Json value() {
        Json j;
        ws();
        if (p >= end) { ok = false; return j; }
        char c = *p;
        if ((c == '{' || c == '[') && ++depth > kMaxDepth) { ok = false; return j; }
        if (c == '{') {
            p++; j.kind = Json::Obj; ws();
            if (p < end && *p == '}') { p++; depth--; return j; }
            while (ok) {
                ws();
                Json k = string_value();
                ws();
                if (p >= end || *p != ':') { ok = false; break; }
                p++;
                Json v = value();
                j.obj.emplace_back(k.str, std::move(v));
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; break; }
                ok = false;
            }
            depth--;
        } else if (c == '[') {
            p++; j.kind = Json::Arr; ws();
            if (p < end && *p == ']') { p++; depth--; return j; }
            while (ok) {
                j.arr.push_back(value());
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; break; }
                ok = false;
            }
            depth--;
        } else if (c == '"') {
            j = string_value();
        } else if (lit("true")) { j.kind = Json::Bool; j.b = true; }
        else if (lit("false")) { j.kind = Json::Bool; j.b = false; }
        else if (lit("null")) { j.kind = Json::Null; }
        else {
            char *e = nullptr;
            j.kind = Json::Num;
            j.num = std::strtod(p, &e);
            if (e == p) ok = false;
            p = e;
        }
        return j;
    }
    Json string_value() {
        Json j; j.kind = Json::Str;
        if (p >= end || *p != '"') { ok = false; return j; }
        p++;
        while (p < end && *p != '"') {
            if (*p == '\\' && p + 1 < end) {
                p++;
                switch (*p) {
                    case 'n': j.str += '\n'; break;
                    case 't': j.str += '\t'; break;
                    case 'r': j.str += '\r'; break;
                    case 'b': j.str += '\b'; break;
                    case 'f': j.str += '\f'; break;
                    case 'u': j.str += '?'; p += (end - p > 4) ? 4 : (end - p - 1); break;     // names only; not needed for geometry
                    default: j.str += *p; break;
                }
                p++;
            } else j.str += *p++;
        }
        if (p >= end) { ok = false; return j; }
        p++;
        return j;
    }
};

struct Accessor {
    const uint8_t *base = nullptr;
    size_t stride = 0, count = 0;
    int component = 0, ncomp = 0;
    bool normalized = false;
};

int component_size(int ct) {
    switch (ct) { case 5120: case 5121: return 1; case 5122: case 5123: return 2; case 5125: case 5126: return 4; }
    return 0;
}
int type_components(const std::string &t) {
    static const struct { const char *name; int n; } table[] = {
        {"SCALAR", 1}, {"VEC2", 2}, {"VEC3", 3}, {"VEC4", 4}, {"MAT2", 4}, {"MAT3", 9}, {"MAT4", 16}};
    for (auto &e : table) if (t == e.name) return e.n;
    return 0;
}

bool resolve_accessor(const Json &root, long index, const std::vector<uint8_t> &bin, Accessor &out) {
    const Json *accs = root.get("accessors");
    if (!accs || index < 0 || (size_t)index >= accs->size()) return false;
    const Json &a = accs->arr[index];
    const Json *bvi = a.get("bufferView");
    if (!bvi) return false;                                      // sparse / zero-filled accessors: not supported
    const Json *views = root.get("bufferViews");
    if (!views || bvi->as_int(-1) < 0 || (size_t)bvi->as_int(-1) >= views->size()) return false;
    const Json &v = views->arr[bvi->as_int(0)];
    if (v.get("buffer") && v.get("buffer")->as_int(0) != 0) return false;   // GLB: buffer 0 is the BIN chunk
    // every integer read from the file is range-checked before it becomes a size: a negative or huge value must not wrap
    auto nonneg = [](const Json *j, long dflt, size_t &out) { long x = j ? j->as_int(dflt) : dflt; if (x < 0) return false; out = (size_t)x; return true; };
    size_t voff = 0, aoff = 0, bstride = 0;
    if (!nonneg(v.get("byteOffset"), 0, voff) || !nonneg(a.get("byteOffset"), 0, aoff) || !nonneg(a.get("count"), 0, out.count) ||
        !nonneg(v.get("byteStride"), 0, bstride))
        return false;
    out.component = (int)(a.get("componentType") ? a.get("componentType")->as_int(0) : 0);
    out.ncomp = a.get("type") ? type_components(a.get("type")->str) : 0;
    out.normalized = a.get("normalized") && a.get("normalized")->b;
    size_t elem = (size_t)component_size(out.component) * out.ncomp;
    if (elem == 0) return false;
    out.stride = bstride ? bstride : elem;
    if (voff > bin.size() || aoff > bin.size() - voff) return false;
    const size_t room = bin.size() - voff - aoff;                            // bytes from the first element to the end of BIN
    if (out.count && (room < elem || (out.count - 1) > (room - elem) / out.stride)) return false;
    out.base = bin.data() + voff + aoff;
    return true;
}

float read_component_float(const uint8_t *p, int ct, bool normalized) {
    switch (ct) {
        case 5126: { float f; std::memcpy(&f, p, 4); return f; }
        case 5120: { int8_t v = (int8_t)*p; return normalized ? std::fmax(v / 127.f, -1.f) : (float)v; }
        case 5121: { uint8_t v = *p; return normalized ? v / 255.f : (float)v; }
        case 5122: { int16_t v; std::memcpy(&v, p, 2); return normalized ? std::fmax(v / 32767.f, -1.f) : (float)v; }
        case 5123: { uint16_t v; std::memcpy(&v, p, 2); return normalized ? v / 65535.f : (float)v; }
        case 5125: { uint32_t v; std::memcpy(&v, p, 4); return (float)v; }
    }
    return 0.f;
}
uint32_t read_index(const uint8_t *p, int ct) {
    switch (ct) {
        case 5121: return *p;
        case 5123: { uint16_t v; std::memcpy(&v, p, 2); return v; }
        case 5125: { uint32_t v; std::memcpy(&v, p, 4); return v; }
    }
    return 0;
}

pt_material fresh_material() {
    pt_material m;
    std::memset(&m, 0, sizeof(m));
    m.alpha = 1.f; m.ior = 1.f; m.roughness = 1.f;      // include/ray.hpp:63-74
    return m;
}

}  // namespace

bool load_glb(const char *path, std::vector<float> &tri_points, std::vector<int32_t> &tri_mats,
              std::vector<pt_material> &materials) {
    FILE *f = std::fopen(path, "rb");
    if (!f) { set_error(std::string("cannot open ") + path); return false; }
    std::vector<uint8_t> file;
    std::fseek(f, 0, SEEK_END);
    long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    file.resize((size_t)sz);
    size_t got = std::fread(file.data(), 1, (size_t)sz, f);
    std::fclose(f);
    if (got != (size_t)sz || sz < 20 || std::memcmp(file.data(), "glTF", 4) != 0) { set_error("not a GLB file"); return false; }
    uint32_t version, total;
    std::memcpy(&version, &file[4], 4);
    std::memcpy(&total, &file[8], 4);
    if (version != 2 || total > file.size()) { set_error("unsupported GLB version/length"); return false; }

    std::string json_text;
    std::vector<uint8_t> bin;
    size_t pos = 12;
    while (pos + 8 <= total) {
        uint32_t clen, ctype;
        std::memcpy(&clen, &file[pos], 4);
        std::memcpy(&ctype, &file[pos + 4], 4);
        if (pos + 8 + (size_t)clen > total) { set_error("truncated GLB chunk"); return false; }
        if (ctype == 0x4E4F534Au) json_text.assign((const char *)&file[pos + 8], clen);            // "JSON"
        else if (ctype == 0x004E4942u && bin.empty()) bin.assign(&file[pos + 8], &file[pos + 8] + clen);   // "BIN\0"
        pos += 8 + (size_t)clen;
    }
    Parser ps{json_text.data(), json_text.data() + json_text.size()};
    Json root = ps.value();
    if (!ps.ok || root.kind != Json::Obj) { set_error("GLB JSON chunk does not parse"); return false; }

    const Json *scenes = root.get("scenes"), *nodes = root.get("nodes"), *meshes = root.get("meshes");
    const Json *mats = root.get("materials");
    if (!scenes || !nodes) { set_error("GLB has no scenes/nodes"); return false; }

    std::vector<long> stack;
    for (auto &scene : scenes->arr)
        if (const Json *roots = scene.get("nodes"))
            for (auto &n : roots->arr) stack.push_back(n.as_int(-1));

    std::vector<char> visited(nodes->size(), 0);       // a valid glTF node graph is a forest: a cycle or a shared child is an error
    while (!stack.empty()) {
        long ni = stack.back();
        stack.pop_back();
        if (ni < 0 || (size_t)ni >= nodes->size()) { set_error("GLB node index out of range"); return false; }
        if (visited[ni]) { set_error("GLB node graph is not a forest (a node is reachable twice)"); return false; }
        visited[ni] = 1;
        const Json &node = nodes->arr[ni];
        const Json *mesh_i = node.get("mesh");
        if (mesh_i && meshes && mesh_i->as_int(-1) >= 0 && (size_t)mesh_i->as_int(-1) < meshes->size()) {
            const Json *prims = meshes->arr[mesh_i->as_int(0)].get("primitives");
            for (size_t pi = 0; prims && pi < prims->size(); pi++) {
                const Json &prim = prims->arr[pi];
                long mode = prim.get("mode") ? prim.get("mode")->as_int(4) : 4;
                if (mode != 4) continue;                                  // cgltf_primitive_type_triangles only

                int32_t mat_idx = 1;
                const Json *pm = prim.get("material");
                if (pm && mats && pm->as_int(-1) >= 0 && (size_t)pm->as_int(-1) < mats->size()) {
                    const Json &gm = mats->arr[pm->as_int(0)];
                    if (const Json *pbr = gm.get("pbrMetallicRoughness")) {
                        const Json *bct = pbr->get("baseColorTexture");
                        if (!bct || !bct->get("index")) {                  // texture-less material
                            pt_material m = fresh_material();
                            float bc[3] = {1.f, 1.f, 1.f};
                            if (const Json *fac = pbr->get("baseColorFactor"))
                                for (size_t k = 0; k < 3 && k < fac->size(); k++) bc[k] = (float)fac->arr[k].num;
                            m.albedo = pt_v3{bc[0], bc[1], bc[2]};
                            materials.push_back(m);
                            mat_idx = (int32_t)materials.size() - 1;
                        }
                    }
                }

                const Json *attrs = prim.get("attributes");
                const Json *posi = attrs ? attrs->get("POSITION") : nullptr;
                if (!posi) continue;
                Accessor pa;
                if (!resolve_accessor(root, posi->as_int(-1), bin, pa) || pa.ncomp != 3) { set_error("bad POSITION accessor"); return false; }
                int pcs = component_size(pa.component);
                auto vertex = [&](size_t i, float out[3]) {
                    const uint8_t *p = pa.base + i * pa.stride;
                    for (int k = 0; k < 3; k++) out[k] = read_component_float(p + (size_t)k * pcs, pa.component, pa.normalized);
                };

                if (const Json *ind = prim.get("indices")) {
                    Accessor ia;
                    if (!resolve_accessor(root, ind->as_int(-1), bin, ia) || ia.ncomp != 1) { set_error("bad index accessor"); return false; }
                    if (ia.count % 3 != 0) { set_error("index count is not a multiple of 3"); return false; }
                    for (size_t j = 0; j + 2 < ia.count; j += 3) {
                        for (int c = 0; c < 3; c++) {
                            uint32_t vi = read_index(ia.base + (j + c) * ia.stride, ia.component);
                            if (vi >= pa.count) { set_error("vertex index out of range"); return false; }
                            float v[3]; vertex(vi, v);
                            tri_points.insert(tri_points.end(), v, v + 3);
                        }
                        tri_mats.push_back(mat_idx);
                    }
                } else {
                    for (size_t j = 0; j + 2 < pa.count; j += 3) {
                        for (int c = 0; c < 3; c++) { float v[3]; vertex(j + c, v); tri_points.insert(tri_points.end(), v, v + 3); }
                        tri_mats.push_back(mat_idx);
                    }
                }
            }
        }
        if (const Json *ch = node.get("children"))
            for (auto &c : ch->arr) stack.push_back(c.as_int(-1));
    }
    return true;
}

}  // namespace pth
