// Host half of "frame in": Camera -> FrameUBO, restating blok/include/camera.hpp:25-59 and
// Denoiser::fillFrameUBO (blok/src/renderer_denoising.cpp:655-704) for a first frame without TAA jitter
// (renderer_draw.cpp:52-89). Matrix conventions are glm's: column-major, right-handed lookAt, and the
// -1..1-depth perspective (the reference includes glm before camera.hpp defines GLM_FORCE_DEPTH_ZERO_TO_ONE,
// so that define has no effect — SURVEY.md §8 a8) with proj[1][1] negated.
#include <cmath>
#include <cstring>

#include "../../include/blokb200.h"

namespace {
struct V3 { float x, y, z; };
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline V3 normalize(V3 v) { float s = 1.0f / std::sqrt(dot(v, v)); return {v.x * s, v.y * s, v.z * s}; }
inline float radians(float d) { return d * 0.01745329251994329576923690768489f; }

// m[c*4 + r]
void mat_mul(const float* a, const float* b, float* out) {  // out = a * b, glm order: ((a0*b0 + a1*b1) + a2*b2) + a3*b3
    float r[16];
    for (int c = 0; c < 4; c++)
        for (int row = 0; row < 4; row++)
            r[c * 4 + row] = a[0 * 4 + row] * b[c * 4 + 0] + a[1 * 4 + row] * b[c * 4 + 1] + a[2 * 4 + row] * b[c * 4 + 2] + a[3 * 4 + row] * b[c * 4 + 3];
    std::memcpy(out, r, sizeof r);
}

// General 4x4 inverse by cofactors (the formulation glm::inverse uses: 2x2 sub-determinant products).
void mat_inverse(const float* m, float* out) {
    auto M = [&](int c, int r) { return m[c * 4 + r]; };
    float c00 = M(2, 2) * M(3, 3) - M(3, 2) * M(2, 3), c02 = M(1, 2) * M(3, 3) - M(3, 2) * M(1, 3), c03 = M(1, 2) * M(2, 3) - M(2, 2) * M(1, 3);
    float c04 = M(2, 1) * M(3, 3) - M(3, 1) * M(2, 3), c06 = M(1, 1) * M(3, 3) - M(3, 1) * M(1, 3), c07 = M(1, 1) * M(2, 3) - M(2, 1) * M(1, 3);
    float c08 = M(2, 1) * M(3, 2) - M(3, 1) * M(2, 2), c10 = M(1, 1) * M(3, 2) - M(3, 1) * M(1, 2), c11 = M(1, 1) * M(2, 2) - M(2, 1) * M(1, 2);
    float c12 = M(2, 0) * M(3, 3) - M(3, 0) * M(2, 3), c14 = M(1, 0) * M(3, 3) - M(3, 0) * M(1, 3), c15 = M(1, 0) * M(2, 3) - M(2, 0) * M(1, 3);
    float c16 = M(2, 0) * M(3, 2) - M(3, 0) * M(2, 2), c18 = M(1, 0) * M(3, 2) - M(3, 0) * M(1, 2), c19 = M(1, 0) * M(2, 2) - M(2, 0) * M(1, 2);
    float c20 = M(2, 0) * M(3, 1) - M(3, 0) * M(2, 1), c22 = M(1, 0) * M(3, 1) - M(3, 0) * M(1, 1), c23 = M(1, 0) * M(2, 1) - M(2, 0) * M(1, 1);
    float f0[4] = {c00, c00, c02, c03}, f1[4] = {c04, c04, c06, c07}, f2[4] = {c08, c08, c10, c11};
    float f3[4] = {c12, c12, c14, c15}, f4[4] = {c16, c16, c18, c19}, f5[4] = {c20, c20, c22, c23};
    float v0[4] = {M(1, 0), M(0, 0), M(0, 0), M(0, 0)}, v1[4] = {M(1, 1), M(0, 1), M(0, 1), M(0, 1)};
    float v2[4] = {M(1, 2), M(0, 2), M(0, 2), M(0, 2)}, v3[4] = {M(1, 3), M(0, 3), M(0, 3), M(0, 3)};
    const float sa[4] = {+1, -1, +1, -1}, sb[4] = {-1, +1, -1, +1};
    float inv[16];
    for (int i = 0; i < 4; i++) {
        inv[0 * 4 + i] = (v1[i] * f0[i] - v2[i] * f1[i] + v3[i] * f2[i]) * sa[i];
        inv[1 * 4 + i] = (v0[i] * f0[i] - v2[i] * f3[i] + v3[i] * f4[i]) * sb[i];
        inv[2 * 4 + i] = (v0[i] * f1[i] - v1[i] * f3[i] + v3[i] * f5[i]) * sa[i];
        inv[3 * 4 + i] = (v0[i] * f2[i] - v1[i] * f4[i] + v2[i] * f5[i]) * sb[i];
    }
    float d0 = m[0] * inv[0], d1 = m[1] * inv[4], d2 = m[2] * inv[8], d3 = m[3] * inv[12];
    float det = (d0 + d1) + (d2 + d3);
    float ood = 1.0f / det;
    for (int i = 0; i < 16; i++) out[i] = inv[i] * ood;
}
}  // namespace

namespace {
void camera_matrices(const blokb_camera* cam, uint32_t width, uint32_t height, float* view, float* proj) {
    const float yaw = radians(cam->yaw), pitch = radians(cam->pitch);
    // camera.hpp:25-31 calls the unqualified C `cos` / `sin`, i.e. the double overloads: products in double, one rounding to float
    V3 dir = {float(std::cos(double(yaw)) * std::cos(double(pitch))), float(std::sin(double(pitch))), float(std::sin(double(yaw)) * std::cos(double(pitch)))};
    const V3 fwd = normalize(dir);
    const V3 right = normalize(cross(fwd, V3{0.f, 1.f, 0.f}));      // :34-36
    const V3 up = normalize(cross(right, fwd));                     // :39-41
    const V3 eye = {cam->position[0], cam->position[1], cam->position[2]};
    // glm::lookAtRH(eye, eye + fwd, up)
    const V3 f = normalize(sub(add(eye, fwd), eye));
    const V3 s = normalize(cross(f, up));
    const V3 u = cross(s, f);
    const float v[16] = {s.x, u.x, -f.x, 0.f, s.y, u.y, -f.y, 0.f, s.z, u.z, -f.z, 0.f, -dot(s, eye), -dot(u, eye), dot(f, eye), 1.f};
    std::memcpy(view, v, 64);
    // glm::perspectiveRH_NO(radians(fov), aspect, 0.1, 10000); [1][1] *= -1   (camera.hpp:55-59)
    const float aspect = static_cast<float>(width) / static_cast<float>(height), zn = 0.1f, zf = 10000.0f;
    const float thf = std::tan(radians(cam->fov) / 2.0f);
    std::memset(proj, 0, 64);
    proj[0] = 1.0f / (aspect * thf);
    proj[5] = 1.0f / thf;
    proj[10] = -(zf + zn) / (zf - zn);
    proj[11] = -1.0f;
    proj[14] = -(2.0f * zf * zn) / (zf - zn);
    proj[5] *= -1.0f;
}
float halton(int index, int base) {  // PostProcess::halton (renderer_postprocess.cpp:216-228)
    float result = 0.0f, f = 1.0f / static_cast<float>(base);
    for (int i = index; i > 0; i = i / base) {
        result += f * static_cast<float>(i % base);
        f = f / static_cast<float>(base);
    }
    return result;
}
}  // namespace

extern "C" int blokb_camera_frame_seq(const blokb_camera* cam, const blokb_camera* prevCam, uint32_t width, uint32_t height, uint32_t frameCount,
                                      uint32_t sampleCount, int32_t jitterIndex, blokb_frame_ubo* out) {
    if (!cam || !out || width == 0 || height == 0) return BLOKB_E_ARG;
    std::memset(out, 0, sizeof *out);
    float view[16], base[16], proj[16];
    camera_matrices(cam, width, height, view, base);
    std::memcpy(proj, base, 64);
    if (jitterIndex >= 0) {  // PostProcess::getJitteredProjection (renderer_postprocess.cpp:254-268), sequence :208-214
        const int k = jitterIndex % 16;
        const float jx = halton(k + 1, 2) - 0.5f, jy = halton(k + 1, 3) - 0.5f;
        proj[8] += (2.0f * jx) / static_cast<float>(width);
        proj[9] += (2.0f * jy) / static_cast<float>(height);
        out->jitterOffset[0] = jx; out->jitterOffset[1] = jy;
    }
    std::memcpy(out->view, view, 64);
    std::memcpy(out->proj, proj, 64);
    mat_inverse(view, out->invView);
    mat_inverse(proj, out->invProj);
    if (prevCam) {  // Denoiser::updatePreviousFrameData(view, baseProj, position) of the previous frame (renderer_draw.cpp:314-318)
        float pv[16], pp[16];
        camera_matrices(prevCam, width, height, pv, pp);
        std::memcpy(out->prevView, pv, 64);
        std::memcpy(out->prevProj, pp, 64);
        mat_mul(pp, pv, out->prevViewProj);
        for (int i = 0; i < 3; i++) out->prevCamPos[i] = prevCam->position[i];
    } else {  // hasPreviousFrame == false: prev* = this frame's (jittered) matrices (renderer_denoising.cpp:672-680)
        std::memcpy(out->prevView, view, 64);
        std::memcpy(out->prevProj, proj, 64);
        mat_mul(proj, view, out->prevViewProj);
        for (int i = 0; i < 3; i++) out->prevCamPos[i] = cam->position[i];
    }
    for (int i = 0; i < 3; i++) out->camPos[i] = cam->position[i];
    out->depth = 1;
    out->frameCount = frameCount; out->sampleCount = sampleCount; out->screenWidth = width; out->screenHeight = height;
    // Denoiser::Settings defaults (blok/include/renderer_denoising.hpp:57-76), copied in by fillFrameUBO (renderer_denoising.cpp:686-703)
    out->temporalAlpha = 0.05f; out->momentAlpha = 0.2f; out->varianceClipGamma = 1.5f; out->depthThreshold = 0.1f;
    out->normalThreshold = 0.95f; out->phiColor = 0.5f; out->phiNormal = 128.0f; out->phiDepth = 0.1f;
    out->atrousIteration = 0; out->stepSize = 1; out->varianceBoost = 1.5f; out->minHistoryLength = 4;
    return 0;
}

extern "C" int blokb_camera_frame(const blokb_camera* cam, uint32_t width, uint32_t height, uint32_t frameCount, uint32_t sampleCount,
                                  blokb_frame_ubo* out) {
    return blokb_camera_frame_seq(cam, nullptr, width, height, frameCount, sampleCount, -1, out);
}
