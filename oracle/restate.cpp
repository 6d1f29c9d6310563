// TEST INFRASTRUCTURE ONLY — oracle/_build/libblokoracle.so
//
// Independent CPU restatement of the reference's voxel ray-traversal hot path. It is the checker that
// travels to the GPU box (where /root/reference does not exist); it is pinned in this container, bit for
// bit, against oracle/_ref (the reference's own sources compiled in place) by tests/test_oracle.py::test_restatement_pinned_to_live_reference,
// and against the committed golden vectors in tests/golden/.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
// The product (blok_b200/) never links, imports or calls it.
//
// Follows, function by function (paths relative to the reference root):
//   slab_root / traverse_subchunk   assets/shaders/intersect.rint:47-55, 70-205
//   hit_face                        assets/shaders/intersect.rint:58-68
//   closest_hit                     assets/shaders/hit.rchit:55-76
//   pcg / random_float / init_rng   assets/shaders/raygen.rgen:77-99
//   sample_cosine / sample_ggx      assets/shaders/raygen.rgen:101-130
//   fresnel / sky / motion          assets/shaders/raygen.rgen:132-165
//   raygen_pixel                    assets/shaders/raygen.rgen:167-414
// glm conventions restated where they fix the rounding order: mat4*vec4 = (c0*x + c1*y) + (c2*z + c3*w)
// (external/glm/glm/detail/type_mat4x4.inl:562-573); dot3 = (x*x' + y*y') + z*z'; normalize(v) =
// v * (1/sqrt(dot(v,v))); mix(a,b,t) = a*(1-t) + b*t; reflect(I,N) = I - N*dot(N,I)*2.
// Build: g++ -std=c++20 -O2 -ffp-contract=off (IEEE fp32, no FMA contraction).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct V3 { float x, y, z; };
static inline V3 v3(float a) { return {a, a, a}; }
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
static inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
static inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
static inline float fmin2(float a, float b) { return (b < a) ? b : a; }   // glm::min
static inline float fmax2(float a, float b) { return (a < b) ? b : a; }   // glm::max
static inline V3 vmin(V3 a, V3 b) { return {fmin2(a.x, b.x), fmin2(a.y, b.y), fmin2(a.z, b.z)}; }
static inline V3 vmax(V3 a, V3 b) { return {fmax2(a.x, b.x), fmax2(a.y, b.y), fmax2(a.z, b.z)}; }
static inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
static inline V3 normalize(V3 v) { return v * (1.0f / std::sqrt(dot(v, v))); }
static inline V3 mix(V3 a, V3 b, float t) { return a * (1.0f - t) + b * t; }
static inline float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
static inline V3 reflect(V3 I, V3 N) { return I - N * dot(N, I) * 2.0f; }

struct Node { uint32_t childMask, firstChild, materialId; float occupancy; };           // svo.hpp:23-28
struct Sub { uint32_t nodeOffset, rootNodeIndex, nodeCount, startDepth; V3 worldMin; float size; V3 worldMax; float pad; };  // resources.hpp:170-184
struct Mat { V3 albedo; uint32_t flags; V3 emission; float ior; };                       // material.hpp:88-114
static_assert(sizeof(Node) == 16 && sizeof(Sub) == 48 && sizeof(Mat) == 32, "layout");

struct Counters { uint64_t fetch = 0, aabb = 0, cap = 0, rays = 0; };  // rays = World::trace calls (one per traceRayEXT)

// State of one traceRayEXT: [tmin, tmax] shrinks as hits are accepted (reportIntersectionEXT semantics).
struct RayState {
    V3 o, d; float tmin, tmax;
    bool anyHit, done, hit;
    int32_t prim; uint32_t node, mat, face;
};

// intersect.rint:58-68
static inline uint32_t hit_face(V3 hitPos, V3 center) {
    V3 diff = hitPos - center;
    V3 a = {std::fabs(diff.x), std::fabs(diff.y), std::fabs(diff.z)};
    if (a.x >= a.y && a.x >= a.z) return diff.x > 0.0f ? 0u : 1u;
    if (a.y >= a.z) return diff.y > 0.0f ? 2u : 3u;
    return diff.z > 0.0f ? 4u : 5u;
}

// intersect.rint:70-205 for one (ray, sub-chunk) pair.
static void traverse_subchunk(const Node* nodes, const Sub& sub, int32_t prim, RayState& r, Counters& c) {
    const uint32_t MAX_STACK = 24u, MAX_ITER = 256u;
    V3 sel = {std::fabs(r.d.x) < 1e-6f ? 1e-6f : r.d.x, std::fabs(r.d.y) < 1e-6f ? 1e-6f : r.d.y,
              std::fabs(r.d.z) < 1e-6f ? 1e-6f : r.d.z};                                  // :79
    V3 invDir = v3(1.0f) / sel;
    V3 t0 = (sub.worldMin - r.o) * invDir, t1 = (sub.worldMax - r.o) * invDir;           // :47-55
    V3 lo = vmin(t0, t1), hi = vmax(t0, t1);
    float tNear = fmax2(fmax2(lo.x, lo.y), lo.z), tFar = fmin2(fmin2(hi.x, hi.y), hi.z);
    float tMin = fmax2(tNear, r.tmin), tMax = fmin2(tFar, r.tmax);                        // :85-86
    if (tMin > tMax) return;                                                              // :88
    c.aabb++;
    uint32_t octantMask = 0u;                                                             // :93-97
    if (r.d.x < 0.0f) octantMask |= 1u;
    if (r.d.y < 0.0f) octantMask |= 2u;
    if (r.d.z < 0.0f) octantMask |= 4u;
    struct Item { uint32_t node; V3 center; float half, tEntry, tExit; };
    Item stack[MAX_STACK];
    uint32_t sp = 0u;
    V3 rootSize = sub.worldMax - sub.worldMin;
    stack[sp++] = {sub.nodeOffset + sub.rootNodeIndex, sub.worldMin + rootSize * 0.5f, rootSize.x * 0.5f, tMin, tMax};
    uint32_t iter = 0u;
    while (sp > 0u) {
        if (!(iter++ < MAX_ITER)) { c.cap++; break; }                                     // :123
        Item it = stack[--sp];
        if (it.tEntry >= r.tmax) continue;                                                // :128
        if (it.node >= sub.nodeOffset + sub.nodeCount) continue;                          // :132
        const Node nd = nodes[it.node];                                                   // :133
        c.fetch++;
        if (nd.childMask == 0u) {                                                         // :136-144
            if (nd.occupancy > 0.0f && !r.done) {
                float t = it.tEntry;
                if (t >= r.tmin && t <= r.tmax && !(r.hit && t == r.tmax)) {              // accept: tmin <= t <= tmax
                    r.face = hit_face(r.o + r.d * t, it.center);
                    r.tmax = t; r.hit = true; r.prim = prim; r.node = it.node; r.mat = nd.materialId;
                    if (r.anyHit) r.done = true;
                }
            }
            continue;
        }
        float nextHalf = it.half * 0.5f;
        for (int i = 7; i >= 0; i--) {                                                    // :164-204
            uint32_t ci = uint32_t(i) ^ octantMask;
            if ((nd.childMask & (1u << ci)) == 0u) continue;
            V3 off = {(ci & 1u) ? it.half : -it.half, (ci & 2u) ? it.half : -it.half, (ci & 4u) ? it.half : -it.half};
            V3 cc = it.center + off * 0.5f;
            V3 a = (cc - v3(nextHalf) - r.o) * invDir, b = (cc + v3(nextHalf) - r.o) * invDir;
            V3 mn = vmin(a, b), mx = vmax(a, b);
            float cmin = fmax2(fmax2(mn.x, mn.y), mn.z), cmax = fmin2(fmin2(mx.x, mx.y), mx.z);
            cmin = fmax2(cmin, it.tEntry); cmax = fmin2(cmax, it.tExit);
            if (cmin < cmax && sp < MAX_STACK) stack[sp++] = {sub.nodeOffset + nd.firstChild + ci, cc, nextHalf, cmin, cmax};
        }
    }
}

// Stand-in for the driver BVH (renderer_raytracing.cpp:15-135: one AABB per SubChunkGpu). Candidates are a
// superset of the boxes the ray touches; traverse_subchunk repeats the root slab test itself, so any
// superset gives the reference's result. Slabs of the 16-unit cell lattice are walked front to back along
// the ray's dominant axis (double precision, fattened), stopping once a slab starts beyond the current tmax.
struct World {
    const Node* nodes; const Sub* subs; size_t nSub; const Mat* mats; size_t nMat;
    bool brute = true; double cell = 16.0; int lo[3], dim[3];
    std::vector<int32_t> cellPrim;

    void build(size_t bruteLimit) {
        brute = nSub <= bruteLimit;
        if (brute) return;
        cell = subs[0].size;
        int hi[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
        lo[0] = lo[1] = lo[2] = INT32_MAX;
        auto cellOf = [&](const Sub& s, int a) { const float* m = &s.worldMin.x; return (int)std::llround(double(m[a]) / cell); };
        for (size_t i = 0; i < nSub; i++)
            for (int a = 0; a < 3; a++) { int c = cellOf(subs[i], a); lo[a] = std::min(lo[a], c); hi[a] = std::max(hi[a], c); }
        for (int a = 0; a < 3; a++) dim[a] = hi[a] - lo[a] + 1;
        cellPrim.assign(size_t(dim[0]) * dim[1] * dim[2], -1);
        for (size_t i = 0; i < nSub; i++) {
            int c[3]; for (int a = 0; a < 3; a++) c[a] = cellOf(subs[i], a) - lo[a];
            cellPrim[(size_t(c[2]) * dim[1] + c[1]) * dim[0] + c[0]] = (int32_t)i;
        }
    }

    void trace(RayState& r, Counters& cn) const {
        cn.rays++;
        if (brute) {
            for (size_t i = 0; i < nSub && !r.done; i++) traverse_subchunk(nodes, subs[i], (int32_t)i, r, cn);
            return;
        }
        const float* od = &r.o.x; const float* dd = &r.d.x;
        int a = 0;
        if (std::fabs(dd[1]) > std::fabs(dd[a])) a = 1;
        if (std::fabs(dd[2]) > std::fabs(dd[a])) a = 2;
        const int ax[2] = {(a + 1) % 3, (a + 2) % 3};
        const double da = dd[a], oa = od[a];
        if (da == 0.0) return;
        const double t0 = r.tmin, t1 = r.tmax;
        double p0 = oa + da * t0, p1 = oa + da * t1;
        int i0 = (int)std::floor(std::min(p0, p1) / cell) - 1, i1 = (int)std::floor(std::max(p0, p1) / cell) + 1;
        i0 = std::max(i0, lo[a]); i1 = std::min(i1, lo[a] + dim[a] - 1);
        const int step = da > 0 ? 1 : -1;
        for (int i = (step > 0 ? i0 : i1); i >= i0 && i <= i1 && !r.done; i += step) {
            double ta = (i * cell - oa) / da, tb = ((i + 1) * cell - oa) / da;
            if (ta > tb) std::swap(ta, tb);
            ta = std::max(ta - 1e-3, t0 - 1e-3); tb = std::min(tb + 1e-3, t1 + 1e-3);
            if (ta > tb) continue;
            if (ta > double(r.tmax) + 1e-3) break;   // everything further is behind the committed hit
            int rlo[2], rhi[2];
            for (int k = 0; k < 2; k++) {
                double q0 = od[ax[k]] + double(dd[ax[k]]) * ta, q1 = od[ax[k]] + double(dd[ax[k]]) * tb;
                if (q0 > q1) std::swap(q0, q1);
                rlo[k] = std::max((int)std::floor((q0 - 0.05) / cell), lo[ax[k]]);
                rhi[k] = std::min((int)std::floor((q1 + 0.05) / cell), lo[ax[k]] + dim[ax[k]] - 1);
            }
            for (int jb = rlo[0]; jb <= rhi[0]; jb++)
                for (int jc = rlo[1]; jc <= rhi[1]; jc++) {
                    int cc[3]; cc[a] = i - lo[a]; cc[ax[0]] = jb - lo[ax[0]]; cc[ax[1]] = jc - lo[ax[1]];
                    int32_t p = cellPrim[(size_t(cc[2]) * dim[1] + cc[1]) * dim[0] + cc[0]];
                    if (p >= 0 && !r.done) traverse_subchunk(nodes, subs[p], p, r, cn);
                }
        }
    }
};

static inline RayState make_ray(V3 o, V3 d, float tmin, float tmax, bool anyHit) {
    RayState r; r.o = o; r.d = d; r.tmin = tmin; r.tmax = tmax; r.anyHit = anyHit; r.done = false; r.hit = false;
    r.prim = -1; r.node = r.mat = r.face = 0xFFFFFFFFu; return r;
}

// ---- raygen.rgen -------------------------------------------------------------------------------
struct Frame {  // FrameUBO, std140, 560 bytes (resources.hpp:103-150; raygen.rgen:13-53)
    float view[16], proj[16], invView[16], invProj[16], prevView[16], prevProj[16], prevViewProj[16];
    V3 camPos; float deltaTime; V3 prevCamPos; uint32_t depth;
    uint32_t frameCount, sampleCount, screenWidth, screenHeight;
    float temporalAlpha, momentAlpha, varianceClipGamma, depthThreshold, normalThreshold, phiColor, phiNormal, phiDepth;
    int32_t atrousIteration, stepSize; float varianceBoost; int32_t minHistoryLength;
    float jitter[2], pad[2];
};
static_assert(sizeof(Frame) == 560, "FrameUBO");

struct V4 { float x, y, z, w; };
static inline V4 mat_mul(const float* m, V4 v) {  // column-major; glm association order
    V4 r;
    r.x = (m[0] * v.x + m[4] * v.y) + (m[8] * v.z + m[12] * v.w);
    r.y = (m[1] * v.x + m[5] * v.y) + (m[9] * v.z + m[13] * v.w);
    r.z = (m[2] * v.x + m[6] * v.y) + (m[10] * v.z + m[14] * v.w);
    r.w = (m[3] * v.x + m[7] * v.y) + (m[11] * v.z + m[15] * v.w);
    return r;
}

static const float PI = 3.14159265359f, INV_PI = 0.31830988618f;

static inline uint32_t pcg(uint32_t& s) {                                                // :77-82
    uint32_t old = s; s = old * 747796405u + 2891336453u;
    uint32_t w = ((old >> ((old >> 28u) + 4u)) ^ old) * 277803737u;
    return (w >> 22u) ^ w;
}
static inline float random_float(uint32_t& s) { return float(pcg(s)) / 4294967295.0f; }   // :84-86
static inline uint32_t init_rng(uint32_t px, uint32_t py, uint32_t width, uint32_t frame, uint32_t sample) {  // :92-99
    uint32_t seed = px + py * width; seed ^= frame * 747796405u; seed ^= sample * 1664525u; pcg(seed); pcg(seed); return seed;
}
static inline void basis(V3 N, V3& T, V3& B) {
    V3 up = std::fabs(N.z) < 0.999f ? V3{0.f, 0.f, 1.f} : V3{1.f, 0.f, 0.f};
    T = normalize(cross(up, N)); B = cross(N, T);
}
static inline V3 sample_cosine(float ux, float uy, V3 N) {                                // :101-113
    float r = std::sqrt(ux), phi = 2.0f * PI * uy;
    float x = r * std::cos(phi), y = r * std::sin(phi), z = std::sqrt(fmax2(0.0f, 1.0f - ux));
    V3 T, B; basis(N, T, B);
    return normalize(T * x + B * y + N * z);
}
static inline V3 sample_ggx(float ux, float uy, V3 N, float rough) {                      // :115-130
    float a = rough * rough, a2 = a * a, phi = 2.0f * PI * ux;
    float ct = std::sqrt((1.0f - uy) / (1.0f + (a2 - 1.0f) * uy)), st = std::sqrt(fmax2(0.0f, 1.0f - ct * ct));
    V3 H = {st * std::cos(phi), st * std::sin(phi), ct};
    V3 T, B; basis(N, T, B);
    return normalize(T * H.x + B * H.y + N * H.z);
}
static inline V3 fresnel(float c, V3 F0) { return F0 + (v3(1.0f) - F0) * std::pow(fmax2(1.0f - c, 0.0f), 5.0f); }  // :132-134
static inline V3 sun_dir() { return normalize(V3{0.5f, 0.8f, 0.3f}); }
static inline V3 sky(V3 dir) {                                                            // :136-148
    float t = 0.5f * (dir.y + 1.0f);
    V3 c = mix(V3{0.8f, 0.85f, 0.95f}, V3{0.4f, 0.6f, 0.9f}, t);
    float sd = fmax2(dot(dir, sun_dir()), 0.0f);
    V3 sun = V3{1.0f, 0.95f, 0.8f} * std::pow(sd, 128.0f) * 5.0f, glow = V3{1.0f, 0.9f, 0.7f} * std::pow(sd, 8.0f) * 0.3f;
    return c + sun + glow;
}
static inline bool is_emissive(V3 e) { return dot(e, v3(1.0f)) > 0.01f; }                  // :158-160
static inline float luminance(V3 c) { return dot(c, V3{0.2126f, 0.7152f, 0.0722f}); }      // :163-165

struct Images { float *color, *worldPos, *normalRough, *albedoMetal, *motion; };
static inline void store(float* img, uint32_t W, uint32_t x, uint32_t y, float a, float b, float c, float d) {
    if (!img) return;
    float* p = img + 4 * (size_t(y) * W + x); p[0] = a; p[1] = b; p[2] = c; p[3] = d;
}

// raygen.rgen:167-414 for one pixel. hit.rchit:55-76 is inlined where the payload is filled.
static void raygen_pixel(const World& w, const Frame& f, uint32_t px, uint32_t py, const Images& out, Counters& cn) {
    const float W = float(f.screenWidth), H = float(f.screenHeight);
    const float cux = (float(px) + 0.5f) / W, cuy = (float(py) + 0.5f) / H;                // :170
    V3 fhPos = v3(0.f), fhN = v3(0.f), fhAlb = v3(0.f), fhEm = v3(0.f);
    float fhRough = 0.f, fhMetal = 0.f, fhDepth = 0.f; bool hadFirst = false, fhEmissive = false;
    V3 accum = v3(0.f);
    const V3 sunDir = sun_dir(); const V3 sunRad = {3.0f, 2.9f, 2.7f};

    for (uint32_t s = 0u; s < f.sampleCount; s++) {                                        // :188
        uint32_t rng = init_rng(px, py, f.screenWidth, f.frameCount, s);
        float pcx, pcy;
        if (s == 0u) { pcx = float(px) + 0.5f; pcy = float(py) + 0.5f; }
        else {
            float jx = random_float(rng) - 0.5f; float jy = random_float(rng) - 0.5f;       // :197 (x drawn first)
            pcx = float(px) + 0.5f + jx * 0.5f; pcy = float(py) + 0.5f + jy * 0.5f;
        }
        float dx = (pcx / W) * 2.0f - 1.0f, dy = (pcy / H) * 2.0f - 1.0f;                  // :201-202
        V4 tg = mat_mul(f.invProj, V4{dx, dy, 1.0f, 1.0f});                                // :204
        V3 tn = normalize(V3{tg.x, tg.y, tg.z});
        V4 wd = mat_mul(f.invView, V4{tn.x, tn.y, tn.z, 0.0f});
        V3 rayDir = normalize(V3{wd.x, wd.y, wd.z});                                       // :205
        V3 rayOrigin = f.camPos;
        V3 radiance = v3(0.f), thr = v3(1.f);

        for (uint32_t bounce = 0u; bounce < 2u; bounce++) {                                // :211-212
            RayState r = make_ray(rayOrigin, rayDir, 0.001f, 10000.0f, false);             // :217-229
            w.trace(r, cn);
            if (!r.hit) { radiance = radiance + thr * sky(rayDir); break; }                // :232-235
            // hit.rchit:55-76
            static const V3 FN[6] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
            const Mat& m = w.mats[std::min(r.mat, 65535u)];
            float metallic = float((m.flags >> 24) & 0xFFu) / 255.0f;
            float roughness = fmax2(float((m.flags >> 16) & 0xFFu) / 255.0f, 0.04f);
            V3 N = FN[r.face], albedo = m.albedo, emission = m.emission;
            float hitT = r.tmax;
            V3 hitPos = rayOrigin + rayDir * hitT;                                         // :238
            if (dot(N, rayDir) > 0.0f) N = -N;                                             // :248-250
            if (bounce == 0u && s == 0u && !hadFirst) {                                    // :253-263
                hadFirst = true; fhPos = hitPos; fhN = N; fhAlb = albedo; fhEm = emission; fhRough = roughness;
                fhMetal = metallic; fhDepth = hitT; fhEmissive = is_emissive(emission);
            }
            if (is_emissive(emission)) {                                                   // :265-277
                radiance = radiance + thr * emission;
                if (luminance(emission) > 5.0f || bounce > 0u) break;
            }
            float NdotL = fmax2(dot(N, sunDir), 0.0f);                                      // :280
            if (NdotL > 0.0f && bounce == 0u) {                                            // :282-326
                RayState sr = make_ray(hitPos + N * 0.001f, sunDir, 0.001f, 1000.0f, true);
                w.trace(sr, cn);
                if (!sr.hit) {
                    V3 diffuse = albedo * (1.0f - metallic);
                    radiance = radiance + thr * diffuse * sunRad * NdotL * INV_PI;
                    if (roughness < 0.9f) {
                        V3 Hh = normalize(sunDir - rayDir);
                        float NdotH = fmax2(dot(N, Hh), 0.0f), VdotH = fmax2(dot(-rayDir, Hh), 0.0f);
                        float a = roughness * roughness, a2 = a * a;
                        float denom = NdotH * NdotH * (a2 - 1.0f) + 1.0f;
                        float D = a2 / (PI * denom * denom);
                        V3 F0 = mix(v3(0.04f), albedo, metallic);
                        V3 F = fresnel(VdotH, F0);
                        radiance = radiance + thr * F * D * 0.25f * sunRad * NdotL;
                    }
                }
            }
            if (bounce > 0u) {                                                             // :329-335
                float p = fmin2(fmax2(fmax2(thr.x, thr.y), thr.z), 0.95f);
                if (random_float(rng) > p) break;
                thr = thr / p;
            }
            float ux = random_float(rng); float uy = random_float(rng);                    // :338
            V3 F0 = mix(v3(0.04f), albedo, metallic);
            V3 V = -rayDir;
            float NdotV = fmax2(dot(N, V), 0.001f);
            V3 F = fresnel(NdotV, F0);
            float sw = (F.x + F.y + F.z) / 3.0f;
            sw = mixf(sw, 1.0f, metallic);
            if (random_float(rng) < sw) {                                                  // :349-361
                V3 Hh = sample_ggx(ux, uy, N, fmax2(roughness, 0.04f));
                V3 nd = reflect(rayDir, Hh);
                if (dot(nd, N) <= 0.0f) break;
                float HdotV = fmax2(dot(Hh, V), 0.0f);
                V3 Fh = fresnel(HdotV, F0);
                thr = thr * (Fh / fmax2(sw, 0.001f));
                rayDir = nd;
            } else {                                                                       // :362-367
                V3 nd = sample_cosine(ux, uy, N);
                V3 diffuse = albedo * (1.0f - metallic);
                thr = thr * (diffuse / fmax2(1.0f - sw, 0.001f));
                rayDir = nd;
            }
            float mt = fmax2(fmax2(thr.x, thr.y), thr.z);                                  // :370-373
            if (mt > 10.0f) thr = thr * (10.0f / mt);
            rayOrigin = hitPos + N * 0.002f;                                               // :376
        }
        accum = accum + radiance;
    }
    V3 color = accum / float(f.sampleCount);                                               // :383
    float mv = fmax2(fmax2(color.x, color.y), color.z);
    if (mv > 100.0f) color = color * (100.0f / mv);
    store(out.color, f.screenWidth, px, py, color.x, color.y, color.z, 1.0f);
    if (!hadFirst) {                                                                       // :395-400
        fhDepth = 10000.0f;
        V4 fw = mat_mul(f.invView, V4{0.f, 0.f, -1.f, 0.f});
        fhPos = f.camPos + normalize(V3{fw.x, fw.y, fw.z}) * 10000.0f;
        fhN = V3{0.f, 1.f, 0.f};
        fhAlb = sky(normalize(fhPos - f.camPos));
    }
    V3 finalAlb = fhEmissive ? fhEm : fhAlb;
    store(out.worldPos, f.screenWidth, px, py, fhPos.x, fhPos.y, fhPos.z, fhDepth);
    store(out.normalRough, f.screenWidth, px, py, fhN.x, fhN.y, fhN.z, fhRough);
    store(out.albedoMetal, f.screenWidth, px, py, finalAlb.x, finalAlb.y, finalAlb.z, fhMetal);
    float mx = 0.f, my = 0.f;
    if (hadFirst && fhDepth < 9999.0f) {                                                   // :150-155, :408-411
        V4 pc = mat_mul(f.prevViewProj, V4{fhPos.x, fhPos.y, fhPos.z, 1.0f});
        float nx = pc.x / pc.w, ny = pc.y / pc.w;
        mx = cux - (nx * 0.5f + 0.5f); my = cuy - (ny * 0.5f + 0.5f);
    }
    store(out.motion, f.screenWidth, px, py, mx, my, 0.f, 0.f);
}

template <class F>
static void parallel_for(size_t n, int nThreads, F&& fn) {
    if (nThreads <= 1) { fn(0, n); return; }
    std::vector<std::thread> th; std::atomic<size_t> next{0}; const size_t grain = 4096;
    for (int t = 0; t < nThreads; t++)
        th.emplace_back([&] { for (;;) { size_t b = next.fetch_add(grain); if (b >= n) break; fn(b, std::min(n, b + grain)); } });
    for (auto& x : th) x.join();
}
}  // namespace

extern "C" {
// Same call shapes as oracle/ref_harness.cpp's ref_trace / ref_render.
int orc_trace(const void* nodes, uint64_t, const void* subs, uint64_t nSub, const float* rays, uint64_t n, int anyHit,
              uint32_t* outHit, float* outT, uint64_t* counters, int nThreads) {
    World w{(const Node*)nodes, (const Sub*)subs, (size_t)nSub, nullptr, 0}; w.build(64);
    std::atomic<uint64_t> c0{0}, c1{0}, c2{0}, c3{0};
    parallel_for(n, nThreads, [&](size_t b, size_t e) {
        Counters cn;
        for (size_t i = b; i < e; i++) {
            const float* q = rays + 8 * i;
            RayState r = make_ray({q[0], q[1], q[2]}, {q[3], q[4], q[5]}, q[6], q[7], anyHit != 0);
            w.trace(r, cn);
            uint32_t* o = outHit + 4 * i;
            if (r.hit) { o[0] = (uint32_t)r.prim; o[1] = r.node; o[2] = r.mat; o[3] = r.face; outT[i] = r.tmax; }
            else { o[0] = o[1] = o[2] = o[3] = 0xFFFFFFFFu; outT[i] = -1.0f; }
        }
        c0 += cn.fetch; c1 += cn.aabb; c2 += cn.cap; c3 += cn.rays;
    });
    if (counters) { counters[0] += c0; counters[1] += c1; counters[2] += c2; counters[3] += c3; }
    return 0;
}

int orc_render(const void* nodes, uint64_t, const void* subs, uint64_t nSub, const void* mats, uint64_t nMat,
               const void* ubo560, float* color, float* worldPos, float* normalRough, float* albedoMetal, float* motion,
               uint64_t pixBegin, uint64_t pixEnd, uint64_t pixStep, uint64_t* counters, int nThreads) {
    World w{(const Node*)nodes, (const Sub*)subs, (size_t)nSub, (const Mat*)mats, (size_t)nMat}; w.build(64);
    Frame f; std::memcpy(&f, ubo560, 560);
    Images im{color, worldPos, normalRough, albedoMetal, motion};
    const uint64_t total = uint64_t(f.screenWidth) * f.screenHeight;
    if (pixEnd > total) pixEnd = total;
    const uint64_t cnt = pixEnd > pixBegin ? (pixEnd - pixBegin + pixStep - 1) / pixStep : 0;
    std::atomic<uint64_t> c0{0}, c1{0}, c2{0}, c3{0};
    parallel_for(cnt, nThreads, [&](size_t b, size_t e) {
        Counters cn;
        for (size_t i = b; i < e; i++) {
            uint64_t p = pixBegin + i * pixStep;
            raygen_pixel(w, f, uint32_t(p % f.screenWidth), uint32_t(p / f.screenWidth), im, cn);
        }
        c0 += cn.fetch; c1 += cn.aabb; c2 += cn.cap; c3 += cn.rays;
    });
    if (counters) { counters[0] += c0; counters[1] += c1; counters[2] += c2; counters[3] += c3; }
    return 0;
}

// Primary rays of sample 0 (pixel centre), raygen.rgen:193-206. rays: w*h x {o,d,tmin=0.001,tmax=10000}.
void orc_primary_rays(const void* ubo560, float* rays) {
    Frame f; std::memcpy(&f, ubo560, 560);
    const float W = float(f.screenWidth), H = float(f.screenHeight);
    for (uint32_t y = 0; y < f.screenHeight; y++)
        for (uint32_t x = 0; x < f.screenWidth; x++) {
            float dx = ((float(x) + 0.5f) / W) * 2.0f - 1.0f, dy = ((float(y) + 0.5f) / H) * 2.0f - 1.0f;
            V4 tg = mat_mul(f.invProj, V4{dx, dy, 1.0f, 1.0f});
            V3 tn = normalize(V3{tg.x, tg.y, tg.z});
            V4 wd = mat_mul(f.invView, V4{tn.x, tn.y, tn.z, 0.0f});
            V3 d = normalize(V3{wd.x, wd.y, wd.z});
            float* r = rays + 8 * (size_t(y) * f.screenWidth + x);
            r[0] = f.camPos.x; r[1] = f.camPos.y; r[2] = f.camPos.z; r[3] = d.x; r[4] = d.y; r[5] = d.z; r[6] = 0.001f; r[7] = 10000.0f;
        }
}

// FNV-1a-64 over bytes (golden-vector checksum).
uint64_t orc_fnv1a64(const void* data, uint64_t nBytes) {
    const uint8_t* p = (const uint8_t*)data; uint64_t h = 0xCBF29CE484222325ull;
    for (uint64_t i = 0; i < nBytes; i++) { h ^= p[i]; h *= 0x100000001B3ull; }
    return h;
}
int orc_hardware_threads() { return (int)std::thread::hardware_concurrency(); }
}
