// TEST INFRASTRUCTURE ONLY — oracle/_ref/libblokref.so
//
// CPU harness around the *reference's own sources*, compiled where they lie under $BLOK_REF:
//   * scene build: blok/src/{svo,material,chunk_manager,vox_loader}.cpp, unmodified, following the
//     recipe of blok/src/app.cpp:100-121 (ChunkManager(128,1) -> loadAndImportVox -> rebuildDirtyChunks
//     -> packChunksToGpuSvo);
//   * traversal + shading: assets/shaders/{intersect.rint, hit.rchit, raygen.rgen}, rewritten to C++ at
//     build time by oracle/glsl2cpp.py into oracle/_ref/gen/ (never committed) and #included below.
// This file supplies only what the Vulkan driver supplies in the reference: the gl_* built-ins, the
// candidate loop over sub-chunk AABBs (renderer_raytracing.cpp:15-30 builds one AABB per SubChunkGpu),
// reportIntersectionEXT / traceRayEXT semantics, and imageStore.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
#include <glm.hpp>
#include <gtc/matrix_transform.hpp>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "camera.hpp"
#include "chunk_manager.hpp"
#include "material.hpp"
#include "vox_loader.hpp"

// ----------------------------------------------------------------------------------------------
// GLSL environment for the generated fragments
// ----------------------------------------------------------------------------------------------
namespace glslenv {
using namespace glm;
typedef uint32_t uint;

struct accelerationStructureEXT {};
struct image2D {
    float* data = nullptr;  // 4 floats per pixel, row-major
    int width = 0;
};

static const uint gl_RayFlagsOpaqueEXT = 1u;
static const uint gl_RayFlagsTerminateOnFirstHitEXT = 4u;
static const uint gl_RayFlagsSkipClosestHitShaderEXT = 8u;

// per-invocation built-ins
static thread_local vec3 gl_WorldRayOriginEXT;
static thread_local vec3 gl_WorldRayDirectionEXT;
static thread_local float gl_RayTminEXT;
static thread_local float gl_RayTmaxEXT;
static thread_local int gl_PrimitiveID;
static thread_local uvec3 gl_LaunchIDEXT;
static thread_local uvec3 gl_LaunchSizeEXT;
static thread_local uint gl_HitKindEXT;
static thread_local float gl_HitTEXT;

// instrumentation written by the (additively) instrumented intersect.rint
static thread_local uint ref_cap_node;
static thread_local uint64_t ref_cnt_fetch, ref_cnt_cap, ref_cnt_aabb, ref_cnt_rays;  // additive counters only: no effect on any result

// committed-hit state of the current traceRayEXT
static thread_local bool cur_any_hit;
static thread_local bool cur_terminated;
static thread_local int hit_prim;
static thread_local uint hit_node, hit_mat, hit_kind;
static thread_local bool hit_valid;

bool reportIntersectionEXT(float t, uint kind);
void traceRayEXT(accelerationStructureEXT&, uint flags, uint cullMask, uint sbtOffset, uint sbtStride,
                 uint missIndex, vec3 origin, float tmin, vec3 dir, float tmax, int payloadLoc);
static inline void imageStore(image2D& img, ivec2 p, vec4 v) {
    if (!img.data) return;
    float* d = img.data + 4 * (size_t(p.y) * img.width + p.x);
    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
}

namespace rint {
#include "gen/intersect.rint.inc"
}
namespace rchit {
#include "gen/hit.rchit.inc"
}
namespace rgen {
#include "gen/raygen.rgen.inc"
}

static_assert(sizeof(rint::SvoNode) == 16 && sizeof(rint::SubChunkGpu) == 48, "layout");
static_assert(sizeof(rchit::MaterialGpu) == 32, "layout");
static_assert(sizeof(rgen::FrameUBO_t) == 560, "FrameUBO std140 mirror must be 560 bytes");
static_assert(sizeof(rgen::RayPayload) == sizeof(rchit::RayPayload), "payload");

// Vulkan semantics: a reported hit is accepted iff tmin <= t <= tmax; accepting shrinks tmax.
bool reportIntersectionEXT(float t, uint kind) {
    if (cur_terminated) return false;
    if (!(t >= gl_RayTminEXT && t <= gl_RayTmaxEXT)) return false;
    // Equal-t hits from two candidates are unspecified in the reference (driver BVH order).
    // Pinned here: the first accepted one is kept when t is equal *and* the candidate index is higher;
    // candidates are visited in ascending prim index, so "lowest (t, prim) wins".
    if (hit_valid && t == gl_RayTmaxEXT) return false;
    gl_RayTmaxEXT = t;
    hit_valid = true;
    hit_prim = gl_PrimitiveID;
    hit_node = ref_cap_node;
    hit_mat = rint::hitAttribs.materialId;
    hit_kind = kind;
    if (cur_any_hit) cur_terminated = true;
    return true;
}
}  // namespace glslenv

// ----------------------------------------------------------------------------------------------
// Stand-in for the driver's BVH: which sub-chunk AABBs get an intersect.rint invocation.
// Any superset of the AABBs the ray touches gives the same result, because intersect.rint repeats the
// root slab test itself (intersect.rint:82-88). Small worlds: every AABB, in index order (exactly what the
// survey probe did). Large worlds: a conservative walk over the 16-unit cell lattice in double precision.
// ----------------------------------------------------------------------------------------------
namespace {
using glslenv::rint::SubChunkGpu;
using glslenv::rint::SvoNode;
using glslenv::rchit::MaterialGpu;

struct TopGrid {
    const SubChunkGpu* subs = nullptr;
    size_t nSub = 0;
    float cell = 16.f;
    int lo[3] = {0, 0, 0}, dim[3] = {0, 0, 0};
    std::vector<int32_t> cellPrim;  // -1 = empty
    bool brute = true;

    void build(const SubChunkGpu* s, size_t n, size_t bruteLimit) {
        subs = s; nSub = n; brute = n <= bruteLimit;
        if (brute || n == 0) return;
        cell = s[0].subChunkSize;
        int hi[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
        lo[0] = lo[1] = lo[2] = INT32_MAX;
        for (size_t i = 0; i < n; i++)
            for (int a = 0; a < 3; a++) {
                int c = (int)std::floor(s[i].worldMin[a] / cell + 0.5f);
                lo[a] = std::min(lo[a], c); hi[a] = std::max(hi[a], c);
            }
        for (int a = 0; a < 3; a++) dim[a] = hi[a] - lo[a] + 1;
        cellPrim.assign(size_t(dim[0]) * dim[1] * dim[2], -1);
        for (size_t i = 0; i < n; i++) {
            int c[3];
            for (int a = 0; a < 3; a++) c[a] = (int)std::floor(s[i].worldMin[a] / cell + 0.5f) - lo[a];
            cellPrim[(size_t(c[2]) * dim[1] + c[1]) * dim[0] + c[0]] = (int32_t)i;
        }
    }

    void candidates(const glm::vec3& o, const glm::vec3& d, float tmin, float tmax, std::vector<int32_t>& out) const {
        out.clear();
        if (brute) { for (size_t i = 0; i < nSub; i++) out.push_back((int32_t)i); return; }
        int a = 0;
        if (std::fabs(d[1]) > std::fabs(d[a])) a = 1;
        if (std::fabs(d[2]) > std::fabs(d[a])) a = 2;
        const int b = (a + 1) % 3, c = (a + 2) % 3;
        const double da = d[a], oa = o[a], cs = cell;
        if (da == 0.0) return;
        double p0 = oa + da * double(tmin), p1 = oa + da * double(tmax);
        if (p0 > p1) std::swap(p0, p1);
        int i0 = std::max((int)std::floor(p0 / cs) - 1, lo[a]);
        int i1 = std::min((int)std::floor(p1 / cs) + 1, lo[a] + dim[a] - 1);
        for (int i = i0; i <= i1; i++) {
            double ta = (i * cs - oa) / da, tb = ((i + 1) * cs - oa) / da;
            if (ta > tb) std::swap(ta, tb);
            ta = std::max(ta - 1e-3, double(tmin) - 1e-3);
            tb = std::min(tb + 1e-3, double(tmax) + 1e-3);
            if (ta > tb) continue;
            int rlo[2], rhi[2];
            const int ax[2] = {b, c};
            for (int k = 0; k < 2; k++) {
                double q0 = o[ax[k]] + double(d[ax[k]]) * ta, q1 = o[ax[k]] + double(d[ax[k]]) * tb;
                if (q0 > q1) std::swap(q0, q1);
                rlo[k] = std::max((int)std::floor((q0 - 0.05) / cs), lo[ax[k]]);
                rhi[k] = std::min((int)std::floor((q1 + 0.05) / cs), lo[ax[k]] + dim[ax[k]] - 1);
            }
            for (int jb = rlo[0]; jb <= rhi[0]; jb++)
                for (int jc = rlo[1]; jc <= rhi[1]; jc++) {
                    int cc[3];
                    cc[a] = i - lo[a]; cc[b] = jb - lo[b]; cc[c] = jc - lo[c];
                    int32_t p = cellPrim[(size_t(cc[2]) * dim[1] + cc[1]) * dim[0] + cc[0]];
                    if (p >= 0) out.push_back(p);
                }
        }
        std::sort(out.begin(), out.end());
        out.erase(std::unique(out.begin(), out.end()), out.end());
    }
};

struct World {
    const SvoNode* nodes = nullptr; size_t nNodes = 0;
    const SubChunkGpu* subs = nullptr; size_t nSub = 0;
    const MaterialGpu* mats = nullptr; size_t nMat = 0;
    TopGrid grid;
};
static const World* g_world = nullptr;  // bound for the duration of one exported call
static thread_local std::vector<int32_t> tl_cand;

struct HitRec { int32_t prim; uint32_t node, mat, face; float t; };

// One traceRayEXT: run intersect.rint for each candidate AABB; closest accepted report wins.
static bool trace_one(const glm::vec3& o, const glm::vec3& d, float tmin, float tmax, bool anyHit, HitRec* rec) {
    using namespace glslenv;
    gl_WorldRayOriginEXT = o; gl_WorldRayDirectionEXT = d;
    gl_RayTminEXT = tmin; gl_RayTmaxEXT = tmax;
    cur_any_hit = anyHit; cur_terminated = false; hit_valid = false;
    ref_cnt_rays++;
    g_world->grid.candidates(o, d, tmin, tmax, tl_cand);
    for (int32_t p : tl_cand) {
        gl_PrimitiveID = p;
        rint::shader_main();
        if (cur_terminated) break;
    }
    if (rec && hit_valid) { rec->prim = hit_prim; rec->node = hit_node; rec->mat = hit_mat; rec->face = hit_kind; rec->t = gl_RayTmaxEXT; }
    return hit_valid;
}
}  // namespace

namespace glslenv {
// raygen.rgen:217-229 (closest hit -> hit.rchit) and :286-298 (shadow: terminate on first hit, skip
// closest-hit shader, miss shader index 1 = shadow.rmiss which clears isShadowed).
void traceRayEXT(accelerationStructureEXT&, uint flags, uint, uint, uint, uint missIndex, vec3 origin, float tmin,
                 vec3 dir, float tmax, int payloadLoc) {
    const bool anyHit = (flags & gl_RayFlagsTerminateOnFirstHitEXT) != 0;
    const bool skipChit = (flags & gl_RayFlagsSkipClosestHitShaderEXT) != 0;
    HitRec rec{};
    const bool hit = trace_one(origin, dir, tmin, tmax, anyHit, &rec);
    if (hit) {
        if (!skipChit) {
            gl_HitKindEXT = rec.face; gl_HitTEXT = rec.t;
            rint::hitAttribs.materialId = rec.mat;
            rchit::hitAttribs.materialId = rec.mat;
            std::memcpy(&rchit::payload, &rgen::payload, sizeof(rgen::payload));
            rchit::shader_main();
            std::memcpy(&rgen::payload, &rchit::payload, sizeof(rgen::payload));
        }
    } else {
        // miss.rmiss declares a payload struct that does not match raygen's (miss.rmiss:11-21); its write
        // lands outside raygen's payload. raygen pre-sets hitT = -1 (raygen.rgen:214-215), which is what
        // survives. shadow.rmiss:13-15 clears isShadowed.
        if (missIndex == 1u && payloadLoc == 1) rgen::isShadowed = false;
    }
}
}  // namespace glslenv

// ----------------------------------------------------------------------------------------------
// Exported C API (ctypes-friendly)
// ----------------------------------------------------------------------------------------------
namespace {
struct RefScene {
    blok::WorldSvoGpu world;
    std::vector<blok::MaterialGpu> mats;
    blok::ChunkManager* mgr = nullptr;
    blok::MaterialLibrary* lib = nullptr;
    ~RefScene() { delete mgr; delete lib; }
};

static void finish_scene(RefScene* s) {
    // app.cpp:119-121
    blok::rebuildDirtyChunks(*s->mgr, 1 << 30);
    blok::packChunksToGpuSvo(*s->mgr, s->world);
    s->mats = s->lib->packForGpu();
}

template <class F>
static void parallel_for(size_t n, int nThreads, F&& f) {
    if (nThreads <= 1) { f(0, n, 0); return; }
    std::vector<std::thread> th;
    std::atomic<size_t> next{0};
    const size_t grain = 4096;
    for (int t = 0; t < nThreads; t++)
        th.emplace_back([&, t] {
            for (;;) {
                size_t b = next.fetch_add(grain);
                if (b >= n) break;
                f(b, std::min(n, b + grain), t);
            }
        });
    for (auto& x : th) x.join();
}
}  // namespace

extern "C" {

// Scene build exactly as app.cpp:100-121, with a caller-chosen asset path / offset / model.
void* ref_scene_from_vox(const char* path, float ox, float oy, float oz, uint32_t model) {
    auto* s = new RefScene;
    s->lib = new blok::MaterialLibrary;
    s->mgr = new blok::ChunkManager(128, 1.0f);
    s->mgr->setMaterialLibrary(s->lib);
    std::string err;
    if (!blok::loadAndImportVox(path, *s->mgr, s->lib, glm::vec3(ox, oy, oz), model, &err)) { delete s; return nullptr; }
    finish_scene(s);
    return s;
}

// Same builder fed from a voxel list (ChunkManager::setVoxelMaterial, chunk_manager.cpp:316-328).
// materials: nMat x 32-byte MaterialGpu records are passed through untouched as the material table.
void* ref_scene_from_voxels(const int32_t* xyz, const uint32_t* mat, size_t n, const void* materials, size_t nMat) {
    auto* s = new RefScene;
    s->lib = new blok::MaterialLibrary;
    s->mgr = new blok::ChunkManager(128, 1.0f);
    s->mgr->setMaterialLibrary(s->lib);
    for (size_t i = 0; i < n; i++)
        s->mgr->setVoxelMaterial(glm::vec3((float)xyz[3 * i], (float)xyz[3 * i + 1], (float)xyz[3 * i + 2]), mat[i], 1.0f);
    blok::rebuildDirtyChunks(*s->mgr, 1 << 30);
    blok::packChunksToGpuSvo(*s->mgr, s->world);
    s->mats.resize(nMat);
    if (nMat) std::memcpy(s->mats.data(), materials, nMat * sizeof(blok::MaterialGpu));
    return s;
}

// The same with a density per write (0 erases: the chunk stays in ChunkManager::chunks and is packed as a lone root node).
void* ref_scene_from_voxels_d(const int32_t* xyz, const uint32_t* mat, const float* density, size_t n, const void* materials, size_t nMat) {
    auto* s = new RefScene;
    s->lib = new blok::MaterialLibrary;
    s->mgr = new blok::ChunkManager(128, 1.0f);
    s->mgr->setMaterialLibrary(s->lib);
    for (size_t i = 0; i < n; i++)
        s->mgr->setVoxelMaterial(glm::vec3((float)xyz[3 * i], (float)xyz[3 * i + 1], (float)xyz[3 * i + 2]), mat[i], density[i]);
    blok::rebuildDirtyChunks(*s->mgr, 1 << 30);
    blok::packChunksToGpuSvo(*s->mgr, s->world);
    s->mats.resize(nMat);
    if (nMat) std::memcpy(s->mats.data(), materials, nMat * sizeof(blok::MaterialGpu));
    return s;
}

void ref_scene_counts(void* h, uint64_t* nNodes, uint64_t* nSub, uint64_t* nMat) {
    auto* s = (RefScene*)h;
    *nNodes = s->world.globalNodes.size(); *nSub = s->world.globalSubChunks.size(); *nMat = s->mats.size();
}
void ref_scene_copy(void* h, void* nodes, void* subs, void* mats) {
    auto* s = (RefScene*)h;
    std::memcpy(nodes, s->world.globalNodes.data(), s->world.globalNodes.size() * 16);
    std::memcpy(subs, s->world.globalSubChunks.data(), s->world.globalSubChunks.size() * 48);
    std::memcpy(mats, s->mats.data(), s->mats.size() * 32);
}
// Reference point query (chunk_manager.cpp:330-348): material id of the voxel at integer coords, 0 = empty.
uint32_t ref_scene_voxel_material(void* h, int32_t x, int32_t y, int32_t z) {
    return ((RefScene*)h)->mgr->getVoxelMaterial(glm::vec3((float)x + 0.5f, (float)y + 0.5f, (float)z + 0.5f));
}
void ref_scene_free(void* h) { delete (RefScene*)h; }

// Camera -> FrameUBO exactly as renderer_draw.cpp:52-89 (no TAA jitter, first frame => prev == current):
// view = Camera::view(), proj = Camera::projection(w/h, 0.1, 10000), fillFrameUBO (renderer_denoising.cpp:655-704).
void ref_make_frame(const float* pos, float yaw, float pitch, float fov, uint32_t w, uint32_t h, uint32_t frameCount,
                    uint32_t sampleCount, void* outUbo560) {
    blok::Camera c;
    c.position = glm::vec3(pos[0], pos[1], pos[2]); c.yaw = yaw; c.pitch = pitch; c.fov = fov;
    const float aspect = static_cast<float>(w) / static_cast<float>(h);
    glm::mat4 view = c.view(), proj = c.projection(aspect, 0.1f, 10000.0f);
    glslenv::rgen::FrameUBO_t u{};
    u.view = view; u.proj = proj; u.invView = glm::inverse(view); u.invProj = glm::inverse(proj);
    u.prevView = view; u.prevProj = proj; u.prevViewProj = proj * view;
    u.camPos = c.position; u.deltaTime = 0.f; u.prevCamPos = c.position; u.depth = 1;
    u.frameCount = frameCount; u.sampleCount = sampleCount; u.screenWidth = w; u.screenHeight = h;
    // Denoiser::Settings defaults (blok/include/renderer_denoising.hpp:57-76), copied in by fillFrameUBO (renderer_denoising.cpp:686-703)
    u.temporalAlpha = 0.05f; u.momentAlpha = 0.2f; u.varianceClipGamma = 1.5f; u.depthThreshold = 0.1f;
    u.normalThreshold = 0.95f; u.phiColor = 0.5f; u.phiNormal = 128.0f; u.phiDepth = 0.1f;
    u.atrousIteration = 0; u.stepSize = 1; u.varianceBoost = 1.5f; u.minHistoryLength = 4;
    std::memcpy(outUbo560, &u, 560);
}

// Frame n of a sequence, as Renderer::render builds it (renderer_draw.cpp:52-89, 314-328): jittered projection from
// PostProcess::getJitteredProjection (renderer_postprocess.cpp:254-268; the Halton table of :208-228 is restated here
// because PostProcess itself needs Vulkan), prev* from the previous frame's camera with the non-jittered projection.
static float ref_halton(int index, int base) {
    float result = 0.0f, f = 1.0f / static_cast<float>(base);
    int i = index;
    while (i > 0) { result += f * static_cast<float>(i % base); i = i / base; f = f / static_cast<float>(base); }
    return result;
}
void ref_make_frame_seq(const float* pos, float yaw, float pitch, float fov, const float* prevCam6 /* pos,yaw,pitch,fov or null */, uint32_t w,
                        uint32_t h, uint32_t frameCount, uint32_t sampleCount, int32_t jitterIndex, void* outUbo560) {
    blok::Camera c;
    c.position = glm::vec3(pos[0], pos[1], pos[2]); c.yaw = yaw; c.pitch = pitch; c.fov = fov;
    const float aspect = static_cast<float>(w) / static_cast<float>(h);
    glm::mat4 view = c.view(), baseProj = c.projection(aspect, 0.1f, 10000.0f), proj = baseProj;
    glm::vec2 jitter(0.0f);
    if (jitterIndex >= 0) {
        const int k = jitterIndex % 16;
        jitter = glm::vec2(ref_halton(k + 1, 2) - 0.5f, ref_halton(k + 1, 3) - 0.5f);
        glm::vec2 clip((2.0f * jitter.x) / static_cast<float>(w), (2.0f * jitter.y) / static_cast<float>(h));
        proj[2][0] += clip.x; proj[2][1] += clip.y;
    }
    glslenv::rgen::FrameUBO_t u{};
    u.view = view; u.proj = proj; u.invView = glm::inverse(view); u.invProj = glm::inverse(proj);
    if (prevCam6) {
        blok::Camera p;
        p.position = glm::vec3(prevCam6[0], prevCam6[1], prevCam6[2]); p.yaw = prevCam6[3]; p.pitch = prevCam6[4]; p.fov = prevCam6[5];
        glm::mat4 pv = p.view(), pp = p.projection(aspect, 0.1f, 10000.0f);
        u.prevView = pv; u.prevProj = pp; u.prevViewProj = pp * pv; u.prevCamPos = p.position;
    } else {
        u.prevView = view; u.prevProj = proj; u.prevViewProj = proj * view; u.prevCamPos = c.position;
    }
    u.camPos = c.position; u.deltaTime = 0.f; u.depth = 1;
    u.frameCount = frameCount; u.sampleCount = sampleCount; u.screenWidth = w; u.screenHeight = h;
    u.temporalAlpha = 0.05f; u.momentAlpha = 0.2f; u.varianceClipGamma = 1.5f; u.depthThreshold = 0.1f;
    u.normalThreshold = 0.95f; u.phiColor = 0.5f; u.phiNormal = 128.0f; u.phiDepth = 0.1f;
    u.atrousIteration = 0; u.stepSize = 1; u.varianceBoost = 1.5f; u.minHistoryLength = 4;
    std::memcpy(outUbo560, &u, 560);
    std::memcpy((char*)outUbo560 + 544, &jitter, 8);
}

static void bind_world(World& w, const void* nodes, size_t nNodes, const void* subs, size_t nSub, const void* mats,
                       size_t nMat) {
    w.nodes = (const SvoNode*)nodes; w.nNodes = nNodes; w.subs = (const SubChunkGpu*)subs; w.nSub = nSub;
    w.mats = (const MaterialGpu*)mats; w.nMat = nMat;
    w.grid.build(w.subs, nSub, 64);
    glslenv::rint::nodes = w.nodes; glslenv::rint::subChunks = w.subs; glslenv::rchit::materials = w.mats;
    g_world = &w;
}

// Arbitrary rays through intersect.rint. rays: n x {ox,oy,oz,dx,dy,dz,tmin,tmax}.
// outHit: n x {prim, node, material, face} (0xFFFFFFFF x4 on miss); outT: n (-1 on miss).
// counters[4] += {node fetches (intersect.rint:133), AABB root-test passes (:88), MAX_ITER cap hits, rays traced (traceRayEXT calls)}.
int ref_trace(const void* nodes, uint64_t nNodes, const void* subs, uint64_t nSub, const float* rays, uint64_t n,
              int anyHit, uint32_t* outHit, float* outT, uint64_t* counters, int nThreads) {
    World w; bind_world(w, nodes, nNodes, subs, nSub, nullptr, 0);
    std::atomic<uint64_t> c0{0}, c1{0}, c2{0}, c3{0};
    parallel_for(n, nThreads, [&](size_t b, size_t e, int) {
        glslenv::ref_cnt_fetch = glslenv::ref_cnt_aabb = glslenv::ref_cnt_cap = glslenv::ref_cnt_rays = 0;
        for (size_t i = b; i < e; i++) {
            const float* r = rays + 8 * i;
            HitRec rec{};
            bool hit = trace_one(glm::vec3(r[0], r[1], r[2]), glm::vec3(r[3], r[4], r[5]), r[6], r[7], anyHit != 0, &rec);
            uint32_t* o = outHit + 4 * i;
            if (hit) { o[0] = (uint32_t)rec.prim; o[1] = rec.node; o[2] = rec.mat; o[3] = rec.face; outT[i] = rec.t; }
            else { o[0] = o[1] = o[2] = o[3] = 0xFFFFFFFFu; outT[i] = -1.0f; }
        }
        c0 += glslenv::ref_cnt_fetch; c1 += glslenv::ref_cnt_aabb; c2 += glslenv::ref_cnt_cap; c3 += glslenv::ref_cnt_rays;
    });
    if (counters) { counters[0] += c0; counters[1] += c1; counters[2] += c2; counters[3] += c3; }
    g_world = nullptr;
    return 0;
}

// Full raygen.rgen over pixels [pixBegin, pixEnd) step pixStep of a w x h launch (row-major pixel index).
// Each output image is w*h*4 floats (unquantised values handed to imageStore), may be null.
int ref_render(const void* nodes, uint64_t nNodes, const void* subs, uint64_t nSub, const void* mats, uint64_t nMat,
               const void* ubo560, float* color, float* worldPos, float* normalRough, float* albedoMetal, float* motion,
               uint64_t pixBegin, uint64_t pixEnd, uint64_t pixStep, uint64_t* counters, int nThreads) {
    using namespace glslenv;
    World w; bind_world(w, nodes, nNodes, subs, nSub, mats, nMat);
    std::memcpy(&rgen::frame, ubo560, 560);
    const uint32_t W = rgen::frame.screenWidth, H = rgen::frame.screenHeight;
    rgen::outColor = {color, (int)W}; rgen::outWorldPosition = {worldPos, (int)W};
    rgen::outNormalRoughness = {normalRough, (int)W}; rgen::outAlbedoMetallic = {albedoMetal, (int)W};
    rgen::outMotionVectors = {motion, (int)W};
    if (pixEnd > uint64_t(W) * H) pixEnd = uint64_t(W) * H;
    const uint64_t cnt = pixEnd > pixBegin ? (pixEnd - pixBegin + pixStep - 1) / pixStep : 0;
    std::atomic<uint64_t> c0{0}, c1{0}, c2{0}, c3{0};
    parallel_for(cnt, nThreads, [&](size_t b, size_t e, int) {
        ref_cnt_fetch = ref_cnt_aabb = ref_cnt_cap = ref_cnt_rays = 0;
        for (size_t i = b; i < e; i++) {
            uint64_t p = pixBegin + i * pixStep;
            gl_LaunchIDEXT = uvec3(uint32_t(p % W), uint32_t(p / W), 0u);
            gl_LaunchSizeEXT = uvec3(W, H, 1u);
            rgen::shader_main();
        }
        c0 += ref_cnt_fetch; c1 += ref_cnt_aabb; c2 += ref_cnt_cap; c3 += ref_cnt_rays;
    });
    if (counters) { counters[0] += c0; counters[1] += c1; counters[2] += c2; counters[3] += c3; }
    g_world = nullptr;
    return 0;
}

// Primary rays of sample 0 (pixel centre, no jitter) with the arithmetic of raygen.rgen:193-206, in glm.
// Cross-checked in tests against ref_render's outWorldPosition.w (the hit t of the real raygen).
void ref_primary_rays(const void* ubo560, float* rays /* w*h*8 */) {
    using namespace glm;
    glslenv::rgen::FrameUBO_t f; std::memcpy(&f, ubo560, 560);
    const vec2 pixelSize = vec2(uvec2(f.screenWidth, f.screenHeight));
    for (uint32_t y = 0; y < f.screenHeight; y++)
        for (uint32_t x = 0; x < f.screenWidth; x++) {
            vec2 pixelCenter = vec2(uvec2(x, y)) + vec2(0.5f);
            vec2 uv = pixelCenter / pixelSize;
            vec2 d = uv * 2.0f - 1.0f;
            vec4 target = f.invProj * vec4(d.x, d.y, 1.0f, 1.0f);
            vec3 rayDir = normalize(vec3(f.invView * vec4(normalize(vec3(target)), 0.0f)));
            float* r = rays + 8 * (size_t(y) * f.screenWidth + x);
            r[0] = f.camPos.x; r[1] = f.camPos.y; r[2] = f.camPos.z;
            r[3] = rayDir.x; r[4] = rayDir.y; r[5] = rayDir.z; r[6] = 0.001f; r[7] = 10000.0f;
        }
}

int ref_hardware_threads() { return (int)std::thread::hardware_concurrency(); }
}
