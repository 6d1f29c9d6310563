// MagicaVoxel .vox import: file -> voxel list + material table, i.e. the inputs of
// blokb_upload_voxels, without the reference's ChunkManager / MaterialLibrary in between. Restates, in our own code:
//   loadVoxFile            blok/src/vox_loader.cpp:151-368  (MAIN children: SIZE, XYZI, RGBA, MATL; everything else skipped;
//                                                            a new model starts at each SIZE that follows a non-empty model)
//   VoxFile::getMaterial   blok/src/vox_loader.cpp:116-149  (palette colour + MATL properties -> Material)
//   importVoxMaterials     blok/src/vox_loader.cpp:370-388  (a fresh MaterialLibrary: id 0 = default grey, id i = palette i)
//   MaterialGpu::pack      blok/include/material.hpp:95-111 (flags = metal8<<24 | rough8<<16 | type4<<12 | alpha4<<8 | spec8)
//   importVoxToChunks      blok/src/vox_loader.cpp:390-430  (world = offset + (x, z, y): VOX z is up; voxel = floor(world))
// Pure host code; the file format is Ephtracy's published one (https://paulbourke.net/dataformats/vox/).
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/blokb200.h"

namespace {
thread_local std::string g_voxErr;

struct Reader {
    const uint8_t* p;
    size_t n, pos = 0;
    bool ok = true;
    template <class T> T get() {
        T v{};
        if (pos + sizeof(T) > n) { ok = false; pos = n; return v; }
        std::memcpy(&v, p + pos, sizeof(T));
        pos += sizeof(T);
        return v;
    }
    std::string str() {  // vox_loader.cpp:71-77: length outside 1..1024 reads as ""
        const int32_t len = get<int32_t>();
        if (!ok || len <= 0 || len > 1024) return "";
        std::string s(size_t(len), '\0');
        if (pos + size_t(len) > n) { ok = false; pos = n; return s; }
        std::memcpy(s.data(), p + pos, size_t(len));
        pos += size_t(len);
        return s;
    }
};

struct VoxMat {  // VoxMaterial defaults, blok/include/vox_loader.hpp:27-38
    int type = 0;
    float roughness = 0.5f, metallic = 0.0f, ior = 1.5f, emission = 0.0f, flux = 0.0f, alpha = 1.0f, specular = 0.5f;
    bool has = false;
};
struct Model {
    std::vector<uint8_t> xyzi;  // 4 bytes per voxel
    bool sized = false;
};

float parse_float(const std::string& s, float dflt) {  // std::stof semantics of parseFloat (vox_loader.cpp:108-114)
    try { return std::stof(s); } catch (...) { return dflt; }
}
int parse_type(const std::string& t) { return t == "_metal" ? 1 : t == "_glass" ? 2 : t == "_emit" ? 3 : 0; }
inline float clamp01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }

// MagicaVoxel's default palette: the 6x6x6 cube ff..00 in steps of 0x33 (blue byte fastest, black left out), then ten-step
// red / green / blue / grey ramps. ABGR words; entry 0 is the empty colour.
void default_palette(uint32_t* pal) {
    pal[0] = 0u;
    int k = 1;
    for (int r = 0; r < 6; r++)          // red (low byte) slowest
        for (int g = 0; g < 6; g++)
            for (int b = 0; b < 6; b++) {  // blue (bits 16-23) fastest
                if (b == 5 && g == 5 && r == 5) continue;
                pal[k++] = 0xFF000000u | uint32_t(0xFF - 0x33 * b) << 16 | uint32_t(0xFF - 0x33 * g) << 8 | uint32_t(0xFF - 0x33 * r);
            }
    static const uint32_t ramp[10] = {0xEE, 0xDD, 0xBB, 0xAA, 0x88, 0x77, 0x55, 0x44, 0x22, 0x11};
    for (int c = 0; c < 4; c++)
        for (int i = 0; i < 10; i++) pal[k++] = 0xFF000000u | (c == 3 ? ramp[i] * 0x010101u : ramp[i] << (8 * c));
}
}  // namespace

struct blokb_vox {
    std::vector<int32_t> xyz;
    std::vector<uint32_t> ids;
    std::vector<blokb_material> mats;
    uint32_t nModels = 0;
    uint32_t size[3] = {0, 0, 0};
};

extern "C" {

const char* blokb_vox_last_error(void) { return g_voxErr.c_str(); }

int blokb_vox_parse(const void* data, size_t bytes, uint32_t modelIndex, const float* offset, blokb_vox** out) {
    if (!data || !out) { g_voxErr = "null argument"; return BLOKB_E_ARG; }
    *out = nullptr;
    Reader r{static_cast<const uint8_t*>(data), bytes};
    char magic[4] = {0, 0, 0, 0};
    for (char& c : magic) c = (char)r.get<uint8_t>();
    if (!r.ok || std::memcmp(magic, "VOX ", 4) != 0) { g_voxErr = "not a MagicaVoxel file (the first four bytes are not the \"VOX \" magic)"; return BLOKB_E_ARG; }
    const int32_t version = r.get<int32_t>();
    if (!r.ok) { g_voxErr = "file ends before the format version"; return BLOKB_E_ARG; }
    if (version < 150) { g_voxErr = "format version " + std::to_string(version) + " is older than 150, the oldest this importer reads"; return BLOKB_E_UNSUPPORTED; }
    for (char& c : magic) c = (char)r.get<uint8_t>();
    if (!r.ok || std::memcmp(magic, "MAIN", 4) != 0) { g_voxErr = "the chunk after the header is not MAIN"; return BLOKB_E_ARG; }
    const int32_t mainContent = r.get<int32_t>(), mainChildren = r.get<int32_t>();
    if (!r.ok) { g_voxErr = "file ends inside the MAIN chunk's size fields"; return BLOKB_E_ARG; }
    if (mainContent > 0) r.pos = std::min(r.n, r.pos + size_t(mainContent));
    const size_t end = std::min(r.n, r.pos + size_t(mainChildren > 0 ? mainChildren : 0));

    uint32_t palette[256];
    default_palette(palette);
    VoxMat vm[256];
    std::vector<Model> models;
    std::vector<uint32_t> sizes;  // 3 per model
    Model cur;
    uint32_t curSize[3] = {0, 0, 0};
    while (r.pos < end && r.ok) {
        char id[4];
        for (char& c : id) c = (char)r.get<uint8_t>();
        const int32_t content = r.get<int32_t>(), children = r.get<int32_t>();
        if (!r.ok) break;
        const size_t chunkEnd = r.pos + size_t(content > 0 ? content : 0);
        if (!std::memcmp(id, "SIZE", 4)) {
            if (cur.sized && !cur.xyzi.empty()) {
                models.push_back(std::move(cur)); sizes.insert(sizes.end(), curSize, curSize + 3);
                cur = Model{};
            }
            for (auto& s : curSize) s = (uint32_t)r.get<int32_t>();
            cur.sized = true;
        } else if (!std::memcmp(id, "XYZI", 4)) {
            const int32_t nv = r.get<int32_t>();
            if (!r.ok) { g_voxErr = "XYZI chunk ends before its voxel count"; return BLOKB_E_ARG; }
            const size_t avail = r.pos < r.n ? (r.n - r.pos) / 4 : 0;
            const size_t take = std::min<size_t>(nv > 0 ? size_t(nv) : 0, avail);
            cur.xyzi.insert(cur.xyzi.end(), r.p + r.pos, r.p + r.pos + 4 * take);
        } else if (!std::memcmp(id, "RGBA", 4)) {
            for (int i = 0; i < 255; i++) { const uint32_t c = r.get<uint32_t>(); if (r.ok) palette[i + 1] = c; }
        } else if (!std::memcmp(id, "MATL", 4)) {
            const int32_t mid = r.get<int32_t>();
            if (r.ok) {
                std::unordered_map<std::string, std::string> props;
                const int32_t pairs = r.get<int32_t>();
                for (int32_t i = 0; r.ok && i < pairs; i++) {
                    std::string k = r.str(), v = r.str();
                    if (!k.empty()) props[k] = v;
                }
                if (mid >= 0 && mid < 256) {
                    VoxMat& m = vm[mid];
                    m.has = true;
                    auto f = [&](const char* key, float& dst, float dflt) { auto it = props.find(key); if (it != props.end()) dst = parse_float(it->second, dflt); };
                    if (auto it = props.find("_type"); it != props.end()) m.type = parse_type(it->second);
                    f("_rough", m.roughness, 0.5f); f("_metal", m.metallic, 0.0f); f("_ior", m.ior, 1.5f); f("_emit", m.emission, 0.0f);
                    f("_flux", m.flux, 0.0f); f("_alpha", m.alpha, 1.0f); f("_sp", m.specular, 0.5f);
                }
            }
            r.ok = true;  // a truncated dictionary is not fatal in the reference either: it seeks to the chunk end
        }
        r.pos = std::min(r.n, chunkEnd);
        if (children > 0) r.pos = std::min(r.n, r.pos + size_t(children));
    }
    if (cur.sized || !cur.xyzi.empty()) { models.push_back(std::move(cur)); sizes.insert(sizes.end(), curSize, curSize + 3); }
    if (models.empty()) { g_voxErr = "the file holds no model (no SIZE + XYZI pair with voxels)"; return BLOKB_E_ARG; }
    if (modelIndex >= models.size()) { g_voxErr = "model " + std::to_string(modelIndex) + " requested, but the file has " + std::to_string(models.size()); return BLOKB_E_ARG; }

    blokb_vox* v = new (std::nothrow) blokb_vox;
    if (!v) { g_voxErr = "out of host memory"; return BLOKB_E_ARG; }
    v->nModels = (uint32_t)models.size();
    for (int a = 0; a < 3; a++) v->size[a] = sizes[3 * modelIndex + a];
    // materials: id 0 = MaterialLibrary's default, id i = palette entry i
    v->mats.resize(256);
    v->mats[0] = blokb_material{{0.8f, 0.8f, 0.8f}, (0u << 24) | (127u << 16) | (0u << 12) | (15u << 8) | 127u, {0.f, 0.f, 0.f}, 0.0f};
    for (int i = 1; i < 256; i++) {
        const uint32_t c = palette[i];
        const float albedo[3] = {float(c & 255u) / 255.0f, float((c >> 8) & 255u) / 255.0f, float((c >> 16) & 255u) / 255.0f};
        float alpha = float(c >> 24) / 255.0f, metallic = 0.0f, roughness = 0.5f, ior = 1.5f, specular = 0.5f, emissionPower = 0.0f;
        float emission[3] = {0.f, 0.f, 0.f};
        int type = 0;
        const VoxMat& m = vm[i];
        if (m.has) {
            type = m.type; roughness = m.roughness; metallic = m.metallic; ior = m.ior; specular = m.specular; alpha = m.alpha;
            if (type == 3) {
                for (int a = 0; a < 3; a++) emission[a] = albedo[a];
                emissionPower = m.emission > 0 ? m.emission : m.flux;
                if (emissionPower <= 0) emissionPower = 5.0f;
            }
        }
        blokb_material g;
        for (int a = 0; a < 3; a++) { g.albedo[a] = albedo[a]; g.emission[a] = emission[a] * emissionPower; }
        g.flags = uint32_t(clamp01(metallic) * 255.0f) << 24 | uint32_t(clamp01(roughness) * 255.0f) << 16 | uint32_t(type) << 12 |
                  uint32_t(clamp01(alpha) * 15.0f) << 8 | uint32_t(clamp01(specular) * 255.0f);
        g.ior = type == 2 ? ior : emissionPower;
        v->mats[i] = g;
    }
    const float off[3] = {offset ? offset[0] : 0.f, offset ? offset[1] : 0.f, offset ? offset[2] : 0.f};
    const std::vector<uint8_t>& q = models[modelIndex].xyzi;
    const size_t n = q.size() / 4;
    v->xyz.resize(3 * n);
    v->ids.resize(n);
    for (size_t i = 0; i < n; i++) {
        const float wx = off[0] + float(q[4 * i]), wy = off[1] + float(q[4 * i + 2]), wz = off[2] + float(q[4 * i + 1]);  // VOX z -> y (up)
        v->xyz[3 * i] = (int32_t)std::floor(wx); v->xyz[3 * i + 1] = (int32_t)std::floor(wy); v->xyz[3 * i + 2] = (int32_t)std::floor(wz);
        v->ids[i] = q[4 * i + 3];  // palette index == material id in a fresh library; 0 -> default material
    }
    *out = v;
    return 0;
}

int blokb_vox_load(const char* path, uint32_t modelIndex, const float* offset, blokb_vox** out) {
    if (!path || !out) { g_voxErr = "null argument"; return BLOKB_E_ARG; }
    *out = nullptr;
    FILE* f = std::fopen(path, "rb");
    if (!f) { g_voxErr = std::string("cannot open ") + path; return BLOKB_E_ARG; }
    std::vector<uint8_t> buf;
    uint8_t tmp[1 << 16];
    size_t k;
    while ((k = std::fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + k);
    std::fclose(f);
    return blokb_vox_parse(buf.data(), buf.size(), modelIndex, offset, out);
}
void blokb_vox_counts(const blokb_vox* v, uint64_t* nVoxels, uint64_t* nMaterials, uint32_t* nModels, uint32_t* size3) {
    if (!v) return;
    if (nVoxels) *nVoxels = v->ids.size();
    if (nMaterials) *nMaterials = v->mats.size();
    if (nModels) *nModels = v->nModels;
    if (size3) for (int a = 0; a < 3; a++) size3[a] = v->size[a];
}
const int32_t* blokb_vox_xyz(const blokb_vox* v) { return v ? v->xyz.data() : nullptr; }
const uint32_t* blokb_vox_material_ids(const blokb_vox* v) { return v ? v->ids.data() : nullptr; }
const blokb_material* blokb_vox_materials(const blokb_vox* v) { return v ? v->mats.data() : nullptr; }
void blokb_vox_free(blokb_vox* v) { delete v; }
}
