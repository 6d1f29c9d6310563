// Ray generation and shading arithmetic of the reference's raygen.rgen / hit.rchit, as device functions.
// Operation order follows the shader source statement by statement (and glm's association order where the
// shader calls a glm-equivalent built-in), and the file is compiled with --fmad=false, so +,-,*,/ and sqrt
// round exactly like the CPU oracle; sinf/cosf/powf are CUDA's (≤ 2 ulp from glibc's) — radiance parity is
// therefore a tolerance, hit IDs are exact.
#pragma once
#include <cstdint>

#include "kernels.h"
#include "traverse.cuh"

namespace blokb {

struct F3 { float x, y, z; };
BLOKB_HD F3 f3(float a, float b, float c) { return {a, b, c}; }
BLOKB_HD F3 f3(float a) { return {a, a, a}; }
BLOKB_HD F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
BLOKB_HD F3 operator-(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
BLOKB_HD F3 operator*(F3 a, F3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
BLOKB_HD F3 operator/(F3 a, F3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
BLOKB_HD F3 operator*(F3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
BLOKB_HD F3 operator/(F3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
BLOKB_HD F3 operator-(F3 a) { return {-a.x, -a.y, -a.z}; }
BLOKB_HD float dot3(F3 a, F3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
BLOKB_HD F3 cross3(F3 a, F3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
BLOKB_HD F3 normalize3(F3 v) { return v * (1.0f / sqrtf(dot3(v, v))); }
BLOKB_HD F3 mix3(F3 a, F3 b, float t) { return a * (1.0f - t) + b * t; }
BLOKB_HD F3 reflect3(F3 I, F3 N) { return I - N * dot3(N, I) * 2.0f; }
BLOKB_HD float max3(F3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }

struct F4 { float x, y, z, w; };
BLOKB_HD F4 mat_mul(const float* m, float x, float y, float z, float w) {  // column-major, glm order
    F4 r;
    r.x = (m[0] * x + m[4] * y) + (m[8] * z + m[12] * w);
    r.y = (m[1] * x + m[5] * y) + (m[9] * z + m[13] * w);
    r.z = (m[2] * x + m[6] * y) + (m[10] * z + m[14] * w);
    r.w = (m[3] * x + m[7] * y) + (m[11] * z + m[15] * w);
    return r;
}

#define BLOKB_PI 3.14159265359f
#define BLOKB_INV_PI 0.31830988618f

BLOKB_HD uint32_t pcg(uint32_t& s) {  // raygen.rgen:77-82
    const uint32_t old = s;
    s = old * 747796405u + 2891336453u;
    const uint32_t w = ((old >> ((old >> 28u) + 4u)) ^ old) * 277803737u;
    return (w >> 22u) ^ w;
}
BLOKB_HD float random_float(uint32_t& s) { return float(pcg(s)) / 4294967295.0f; }  // :84-86
BLOKB_HD uint32_t init_rng(uint32_t px, uint32_t py, uint32_t width, uint32_t frame, uint32_t sample) {  // :92-99
    uint32_t seed = px + py * width;
    seed ^= frame * 747796405u;
    seed ^= sample * 1664525u;
    pcg(seed);
    pcg(seed);
    return seed;
}
BLOKB_HD void basis(F3 N, F3& T, F3& B) {
    const F3 up = fabsf(N.z) < 0.999f ? f3(0.f, 0.f, 1.f) : f3(1.f, 0.f, 0.f);
    T = normalize3(cross3(up, N));
    B = cross3(N, T);
}
BLOKB_HD F3 sample_cosine(float ux, float uy, F3 N) {  // :101-113
    const float r = sqrtf(ux), phi = 2.0f * BLOKB_PI * uy;
    const float x = r * cosf(phi), y = r * sinf(phi), z = sqrtf(fmaxf(0.0f, 1.0f - ux));
    F3 T, B;
    basis(N, T, B);
    return normalize3(T * x + B * y + N * z);
}
BLOKB_HD F3 sample_ggx(float ux, float uy, F3 N, float rough) {  // :115-130
    const float a = rough * rough, a2 = a * a, phi = 2.0f * BLOKB_PI * ux;
    const float ct = sqrtf((1.0f - uy) / (1.0f + (a2 - 1.0f) * uy)), st = sqrtf(fmaxf(0.0f, 1.0f - ct * ct));
    const F3 H = f3(st * cosf(phi), st * sinf(phi), ct);
    F3 T, B;
    basis(N, T, B);
    return normalize3(T * H.x + B * H.y + N * H.z);
}
// pow(x, n) for the shaders' three constant integer exponents, x in [0, 1]. On the device: repeated multiplication (<= 64 ulp
// from the exact power for n = 128, a few ulp for 5 and 8) instead of the generic powf (~40 instructions, 3.5 % of k_path);
// radiance parity is a 2e-3 tolerance either way (header comment). The host emulation keeps powf, like the oracle.
template <int N> BLOKB_HD float pow_const(float x) {
#if defined(__CUDA_ARCH__)
    static_assert(N == 5 || N == 8 || N == 128, "exponents used by raygen.rgen");
    const float x2 = x * x, x4 = x2 * x2;
    if (N == 5) return x4 * x;
    const float x8 = x4 * x4;
    if (N == 8) return x8;
    const float x16 = x8 * x8, x32 = x16 * x16, x64 = x32 * x32;
    return x64 * x64;
#else
    return powf(x, float(N));
#endif
}
BLOKB_HD F3 fresnel(float c, F3 F0) { return F0 + (f3(1.0f) - F0) * pow_const<5>(fmaxf(1.0f - c, 0.0f)); }  // :132-134
BLOKB_HD F3 sun_dir() { return normalize3(f3(0.5f, 0.8f, 0.3f)); }
BLOKB_HD F3 sky_color(F3 dir) {  // :136-148
    const float t = 0.5f * (dir.y + 1.0f);
    const F3 c = mix3(f3(0.8f, 0.85f, 0.95f), f3(0.4f, 0.6f, 0.9f), t);
    const float sd = fmaxf(dot3(dir, sun_dir()), 0.0f);
    const F3 sun = f3(1.0f, 0.95f, 0.8f) * pow_const<128>(sd) * 5.0f, glow = f3(1.0f, 0.9f, 0.7f) * pow_const<8>(sd) * 0.3f;
    return c + sun + glow;
}
BLOKB_HD bool is_emissive(F3 e) { return dot3(e, f3(1.0f)) > 0.01f; }            // :158-160
BLOKB_HD float luminance(F3 c) { return dot3(c, f3(0.2126f, 0.7152f, 0.0722f)); }  // :163-165

// Primary ray through pixel-space point (pcx, pcy): raygen.rgen:201-206.
BLOKB_HD F3 primary_dir(const FrameDev& f, float pcx, float pcy) {
    const float dx = (pcx / float(f.width)) * 2.0f - 1.0f, dy = (pcy / float(f.height)) * 2.0f - 1.0f;
    const F4 tg = mat_mul(f.invProj, dx, dy, 1.0f, 1.0f);
    const F3 tn = normalize3(f3(tg.x, tg.y, tg.z));
    const F4 wd = mat_mul(f.invView, tn.x, tn.y, tn.z, 0.0f);
    return normalize3(f3(wd.x, wd.y, wd.z));
}

BLOKB_HD F3 face_normal(uint32_t face) {  // hit.rchit:46-53
    const float s = (face & 1u) ? -1.0f : 1.0f;
    const uint32_t a = face >> 1;
    return f3(a == 0 ? s : 0.f, a == 1 ? s : 0.f, a == 2 ? s : 0.f);
}

struct Surface { F3 albedo, emission; float roughness, metallic; };
BLOKB_HD Surface fetch_surface(const DeviceWorld& W, uint32_t materialId) {  // hit.rchit:62-75
    const uint32_t id = materialId < 65535u ? materialId : 65535u;
    // An id beyond the uploaded table is outside the shader's storage buffer: undefined in the reference, an all-zero record
    // under Vulkan's robustBufferAccess. Same here (never a wild read): black, non-emissive, roughness floor.
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (id < W.nMaterials) { a = __ldg(&W.materials[2 * id]); b = __ldg(&W.materials[2 * id + 1]); }
    const uint32_t flags = __float_as_uint(a.w);
    Surface s;
    s.albedo = f3(a.x, a.y, a.z);
    s.emission = f3(b.x, b.y, b.z);
    s.metallic = float((flags >> 24) & 0xFFu) / 255.0f;
    s.roughness = fmaxf(float((flags >> 16) & 0xFFu) / 255.0f, 0.04f);
    return s;
}

// One sample of raygen.rgen's path loop (:188-380) as a loop over the RAYS of the sample — camera ray, sun shadow ray, bounce ray —
// around ONE inlined copy of the walk. The shape is deliberate: with a copy of traverse() per ray kind the compiler finds no
// re-convergence point between them, lanes that skip a ray run ahead into the next copy, and the warp then executes every copy
// once per group of lanes (measured on the c4 frame: 7.1 threads per instruction and 1.7x the time of this form).
// Every expression is the shader's, in the shader's order, so every form adds the same fp32 values in the same order:
//   * iteration bounce == 0: camera ray (:217-235), emission (:265-277), direct light (:280-326), lobe sampling (:337-376). The
//     lobe sampling does not depend on the shadow verdict and draws from the generator after nothing the shadow ray touches, so it
//     is evaluated BEFORE the shadow ray is traced and `emit(br)` hands the bounce ray out at that point (a record in k_path_front,
//     the caller's variables in the fused form); the two direct-light terms are formed before that walk too and only ADDED after it;
//   * iteration bounce == 1 (path_bounce / the inline third pass): only two statements can still change the radiance — the sky
//     term of a miss (:231-235) and the emission of an emissive hit (:265-277); the Russian roulette, lobe choice and next origin
//     after them (:329-376) feed an iteration that never runs (bounce < 2) and a generator that is re-seeded per sample (:92-99).
// The shadow ray starts 0.001 and the bounce ray 0.002 off the voxel the camera ray hit: both begin in the deepest cell of that
// walk's node stack that holds their start voxel (WalkResume, traverse.cuh).
// When `first` is non-null and the primary ray hits, *first receives the G-buffer first-hit data (:253-263).
struct FirstHit { F3 pos, N, albedo, emission; float roughness, metallic, depth; bool had, emissive; };
struct BounceRay { F3 origin, dir, thr; };

// Iteration bounce == 1 given its hit / miss: the term it adds to the radiance (`add` = whether it adds one at all; kept apart from
// "adds zero" so that every form performs exactly the shader's additions).
BLOKB_HD F3 bounce_term(const DeviceWorld& W, const BounceRay& br, bool hit, uint32_t material, bool& add) {
    if (!hit) { add = true; return br.thr * sky_color(br.dir); }
    const Surface m = fetch_surface(W, material);
    add = is_emissive(m.emission);
    return br.thr * m.emission;
}

// kInlineBounce = false: iteration 0 only; `emit(br)` is called once for a path that goes on (k_path_front).
// kInlineBounce = true: the bounce ray is traced as a third pass of the same loop (the fused kernel, k_path); emit is not used.
template <bool kStats, bool kInlineBounce, class Emit>
BLOKB_HD void path_rays(const DeviceWorld& W, const NodeStack& stk, const FrameDev& f, uint32_t px, uint32_t py, uint32_t s,
                        FirstHit* first, uint32_t& rays, uint32_t& steps, F3& radiance, Emit&& emit) {
    uint32_t rng = init_rng(px, py, f.width, f.frameCount, s);
    float pcx, pcy;
    if (s == 0u) { pcx = float(px) + 0.5f; pcy = float(py) + 0.5f; }
    else {
        const float jx = random_float(rng) - 0.5f;
        const float jy = random_float(rng) - 0.5f;
        pcx = float(px) + 0.5f + jx * 0.5f; pcy = float(py) + 0.5f + jy * 0.5f;
    }
    const F3 rayDir = primary_dir(f, pcx, pcy);
    const F3 rayOrigin = f3(f.camPos[0], f.camPos[1], f.camPos[2]);
    radiance = f3(0.f);
    F3 thr = f3(1.f);
    const F3 sunDir = sun_dir();
    const F3 sunRad = f3(3.0f, 2.9f, 2.7f);
    WalkResume resume;
    F3 o = rayOrigin, d = rayDir;
    float tmax = 10000.0f;
    F3 direct = f3(0.f), spec = f3(0.f);
    bool haveSpec = false, wantShadow = false, wantBounce = false;
    BounceRay br;
    int pass = 0;   // 0 camera ray, 1 sun shadow ray, 2 bounce ray
#pragma unroll 1
    for (;;) {
        HitRec h;
        rays++;
        const bool hit = traverse<kStats>(W, stk, o.x, o.y, o.z, d.x, d.y, d.z, 0.001f, tmax, h, steps, /*locate=*/pass > 0, &resume);
        if (pass == 0) {
            if (!hit) {
                radiance = radiance + thr * sky_color(rayDir);
                break;
            }
            const Surface m = fetch_surface(W, h.material);
            F3 N = face_normal(h.facePrim & 7u);
            const F3 hitPos = rayOrigin + rayDir * h.t;
            if (dot3(N, rayDir) > 0.0f) N = -N;
            if (first && s == 0u && !first->had) {
                first->had = true; first->pos = hitPos; first->N = N; first->albedo = m.albedo; first->emission = m.emission;
                first->roughness = m.roughness; first->metallic = m.metallic; first->depth = h.t; first->emissive = is_emissive(m.emission);
            }
            if (is_emissive(m.emission)) {
                radiance = radiance + thr * m.emission;
                if (luminance(m.emission) > 5.0f) break;
            }
            const float NdotL = fmaxf(dot3(N, sunDir), 0.0f);
            wantShadow = NdotL > 0.0f;
            if (wantShadow) {   // direct light (:280-326), everything but the shadow verdict
                const F3 diffuse = m.albedo * (1.0f - m.metallic);
                direct = thr * diffuse * sunRad * NdotL * BLOKB_INV_PI;
                if (m.roughness < 0.9f) {
                    const F3 Hh = normalize3(sunDir - rayDir);
                    const float NdotH = fmaxf(dot3(N, Hh), 0.0f), VdotH = fmaxf(dot3(-rayDir, Hh), 0.0f);
                    const float a = m.roughness * m.roughness, a2 = a * a;
                    const float denom = NdotH * NdotH * (a2 - 1.0f) + 1.0f;
                    const float D = a2 / (BLOKB_PI * denom * denom);
                    const F3 F0 = mix3(f3(0.04f), m.albedo, m.metallic);
                    const F3 F = fresnel(VdotH, F0);
                    spec = thr * F * D * 0.25f * sunRad * NdotL;
                    haveSpec = true;
                }
            }
            {   // the bounce ray (:329-376; no Russian roulette on iteration 0)
                const float ux = random_float(rng);
                const float uy = random_float(rng);
                const F3 F0 = mix3(f3(0.04f), m.albedo, m.metallic);
                const F3 V = -rayDir;
                const float NdotV = fmaxf(dot3(N, V), 0.001f);
                const F3 F = fresnel(NdotV, F0);
                float sw = (F.x + F.y + F.z) / 3.0f;
                sw = sw * (1.0f - m.metallic) + 1.0f * m.metallic;
                F3 nd;
                bool alive = true;
                if (random_float(rng) < sw) {
                    const F3 Hh = sample_ggx(ux, uy, N, fmaxf(m.roughness, 0.04f));
                    nd = reflect3(rayDir, Hh);
                    alive = !(dot3(nd, N) <= 0.0f);
                    const float HdotV = fmaxf(dot3(Hh, V), 0.0f);
                    const F3 Fh = fresnel(HdotV, F0);
                    thr = thr * (Fh / fmaxf(sw, 0.001f));
                } else {
                    nd = sample_cosine(ux, uy, N);
                    const F3 diffuse = m.albedo * (1.0f - m.metallic);
                    thr = thr * (diffuse / fmaxf(1.0f - sw, 0.001f));
                }
                if (alive) {
                    const float mt = max3(thr);
                    if (mt > 10.0f) thr = thr * (10.0f / mt);
                    br.origin = hitPos + N * 0.002f;
                    br.dir = nd;
                    br.thr = thr;
                    if (kInlineBounce) wantBounce = true;
                    else emit(br);
                }
            }
            if (wantShadow) {
                o = hitPos + N * 0.001f;
                d = sunDir;
                tmax = 1000.0f;
                pass = 1;
                continue;
            }
        } else if (pass == 1) {
            if (!hit) {
                radiance = radiance + direct;
                if (haveSpec) radiance = radiance + spec;
            }
        } else {
            bool add;
            const F3 c = bounce_term(W, br, hit, h.material, add);
            if (add) radiance = radiance + c;
            break;
        }
        if (!(kInlineBounce && wantBounce)) break;
        o = br.origin;
        d = br.dir;
        tmax = 10000.0f;
        pass = 2;
    }
}

// Iteration bounce == 1 of a ray that was handed out by path_rays<.., false> (k_bounce: one thread per sorted bounce ray).
template <bool kStats>
BLOKB_HD F3 path_bounce(const DeviceWorld& W, const NodeStack& stk, const BounceRay& br, uint32_t& rays, uint32_t& steps, bool& add) {
    HitRec h;
    rays++;
    // the bounce ray starts on geometry: start voxel located once (Walker::init, locate = true)
    const bool hit = traverse<kStats>(W, stk, br.origin.x, br.origin.y, br.origin.z, br.dir.x, br.dir.y, br.dir.z, 0.001f, 10000.0f, h, steps,
                                      /*locate=*/true);
    return bounce_term(W, br, hit, h.material, add);
}

// The fused form: all of the sample's rays by this thread. The bounce ray gets its own copy of the walk after the two-pass loop:
// measured on the c4 frame, 16 spp at 4K: 45.1 ms this way, 54.7 ms with the bounce ray as a third pass of the loop (80 registers and
// spills under the same occupancy), 69 ms with a copy of the walk per ray kind; the wavefront schedule does the same samples in 37 ms.
template <bool kStats>
BLOKB_HD F3 path_sample(const DeviceWorld& W, const NodeStack& stk, const FrameDev& f, uint32_t px, uint32_t py, uint32_t s,
                        FirstHit* first, uint32_t& rays, uint32_t& steps) {
    F3 radiance;
    BounceRay br;
    bool alive = false;
    path_rays<kStats, false>(W, stk, f, px, py, s, first, rays, steps, radiance, [&](const BounceRay& b) { br = b; alive = true; });
    if (alive) {
        bool add;
        const F3 c = path_bounce<kStats>(W, stk, br, rays, steps, add);
        if (add) radiance = radiance + c;
    }
    return radiance;
}

}  // namespace blokb
