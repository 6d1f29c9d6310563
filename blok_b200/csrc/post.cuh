// Denoiser + post-process arithmetic of the reference's compute shaders, one function per output pixel
// (SURVEY.md §8(f) ranks 1-2):
//   temporal_variance_pixel  temporal_reproject.comp:185-317 fused with variance.comp:101-145 (the variance pass only
//                            needs this pixel's freshly written moments / history length plus the same 3x3
//                            neighbourhood of the raw frame, so it never goes back to memory for them)
//   atrous_pixel             atrous.comp:143-220 (one 5x5 a-trous iteration)
//   taa_tonemap_pixel        taa.comp:116-220 fused with tonemap.comp:109-144 (tonemap reads exactly the texel TAA wrote)
//   sharpen_pixel            sharpen.comp:20-74
// Statement order, association order and glm's min/max/mix/clamp/normalize definitions are followed exactly and the
// file is compiled with --fmad=false, so every +,-,*,/ and sqrt rounds like the CPU checker; only exp() (variance
// depth weight, "neutral" tonemap operator) is CUDA's expf (<= 2 ulp from glibc's). Pinned driver behaviour (half
// rounding, unorm8 conversion, 8-bit bilinear weights) is listed in oracle/ref_post_harness.cpp and DESIGN.md §9.
#pragma once
#include <cstdint>

#include "post.h"
#include "shade.cuh"
#if defined(__CUDACC__)
#include <cuda_fp16.h>
#endif

namespace blokb {
namespace post {

#if defined(__CUDACC__)
BLOKB_D float h2f(uint32_t h) { return __half2float(__ushort_as_half((unsigned short)h)); }
BLOKB_D uint32_t f2h(float f) { return __half_as_ushort(__float2half_rn(f)); }
#else
static inline float h2f(uint32_t h) {
    uint32_t s = (h >> 15) & 1u, e = (h >> 10) & 31u, m = h & 1023u, u;
    if (e == 0) {
        if (m == 0) u = s << 31;
        else { int sh = 0; while (!(m & 1024u)) { m <<= 1; sh++; } u = (s << 31) | uint32_t(113 - sh) << 23 | ((m & 1023u) << 13); }
    } else if (e == 31) u = (s << 31) | 0x7F800000u | (m << 13);
    else u = (s << 31) | ((e + 112u) << 23) | (m << 13);
    return __uint_as_float(u);
}
static inline uint32_t f2h(float f) {
    uint32_t u = __float_as_uint(f), s = (u >> 16) & 0x8000u, a = u & 0x7FFFFFFFu;
    if (a >= 0x7F800000u) return s | 0x7C00u | (a > 0x7F800000u ? 0x200u | ((a >> 13) & 1023u) : 0u);
    if (a >= 0x477FF000u) return s | 0x7C00u;
    if (a < 0x33000001u) return s;
    int e = int(a >> 23) - 127;
    uint32_t m = (a & 0x7FFFFFu) | 0x800000u;
    int shift = e < -14 ? 13 + (-14 - e) : 13;
    uint32_t q = m >> shift, rem = m & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (q & 1u))) q++;
    return s | (e < -14 ? q : (uint32_t(e + 15) << 10) + (q - 1024u));
}
#endif

// glm's definitions (comparison-based, so NaNs propagate exactly as in the checker)
BLOKB_HD float gmin(float a, float b) { return (b < a) ? b : a; }
BLOKB_HD float gmax(float a, float b) { return (a < b) ? b : a; }
BLOKB_HD float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }
BLOKB_HD float gmix(float x, float y, float a) { return x * (1.0f - a) + y * a; }
BLOKB_HD F3 gmin3(F3 a, F3 b) { return {gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z)}; }
BLOKB_HD F3 gmax3(F3 a, F3 b) { return {gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z)}; }
BLOKB_HD F3 abs3(F3 a) { return {fabsf(a.x), fabsf(a.y), fabsf(a.z)}; }
BLOKB_HD F3 sqrt3(F3 a) { return {sqrtf(a.x), sqrtf(a.y), sqrtf(a.z)}; }
BLOKB_HD float length3(F3 a) { return sqrtf(dot3(a, a)); }
BLOKB_HD float lum709(F3 c) { return dot3(c, f3(0.2126f, 0.7152f, 0.0722f)); }
BLOKB_HD F3 operator*(float s, F3 a) { return {s * a.x, s * a.y, s * a.z}; }
BLOKB_HD F3 rgb_to_ycocg(F3 c) {
    return {0.25f * c.x + 0.5f * c.y + 0.25f * c.z, 0.5f * c.x - 0.5f * c.z, -0.25f * c.x + 0.5f * c.y - 0.25f * c.z};
}
BLOKB_HD F3 ycocg_to_rgb(F3 c) { return {c.x + c.y - c.z, c.x + c.z, c.x - c.y - c.z}; }

// ---- image access ------------------------------------------------------------------------------------------
BLOKB_HD int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
BLOKB_HD F4 ld_rgba32f(const float* img, int W, int x, int y) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(img) + (unsigned(y) * unsigned(W) + unsigned(x)));
    return {v.x, v.y, v.z, v.w};
}
BLOKB_HD F4 ld_rgba16f(const uint16_t* img, int W, int x, int y) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(img) + (unsigned(y) * unsigned(W) + unsigned(x)));
    return {h2f(v.x & 0xFFFFu), h2f(v.x >> 16), h2f(v.y & 0xFFFFu), h2f(v.y >> 16)};
}
BLOKB_HD F4 ld_rgba8(const uint8_t* img, int W, int x, int y) {
    const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(img) + (unsigned(y) * unsigned(W) + unsigned(x)));
    return {float(v & 255u) / 255.0f, float((v >> 8) & 255u) / 255.0f, float((v >> 16) & 255u) / 255.0f, float(v >> 24) / 255.0f};
}
BLOKB_HD uint32_t pack_unorm8(float v) {
    float x = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
    if (!(x == x)) x = 0.0f;
    return (uint32_t)floorf(x * 255.0f + 0.5f);
}
BLOKB_HD F4 operator*(F4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
BLOKB_HD F4 operator+(F4 a, F4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }

// The reference's sampler (linear, clamp-to-edge, unnormalised = u*W - 0.5, 8-bit weights); see header comment.
template <bool UNORM8> BLOKB_HD F4 tex_linear(const void* img, int W, int H, float u, float v) {
    const float xb = u * float(W) - 0.5f, yb = v * float(H) - 0.5f;
    const float fx = floorf(xb), fy = floorf(yb);
    int qa = int((xb - fx) * 256.0f + 0.5f), qb = int((yb - fy) * 256.0f + 0.5f);
    const int x0 = !(fx >= 0.0f) ? 0 : (fx > float(W - 1) ? W - 1 : int(fx));
    const int x1 = !(fx + 1.0f >= 0.0f) ? 0 : (fx + 1.0f > float(W - 1) ? W - 1 : int(fx + 1.0f));
    const int y0 = !(fy >= 0.0f) ? 0 : (fy > float(H - 1) ? H - 1 : int(fy));
    const int y1 = !(fy + 1.0f >= 0.0f) ? 0 : (fy + 1.0f > float(H - 1) ? H - 1 : int(fy + 1.0f));
    if (!(xb == xb)) qa = 0;
    if (!(yb == yb)) qb = 0;
    auto ld = [&](int x, int y) {
        if constexpr (UNORM8) return ld_rgba8(static_cast<const uint8_t*>(img), W, x, y);
        else return ld_rgba32f(static_cast<const float*>(img), W, x, y);
    };
    if ((qa == 0 || qa == 256) && (qb == 0 || qb == 256)) return ld(qa ? x1 : x0, qb ? y1 : y0);  // texel-exact fast path
    const float a = float(qa) / 256.0f, b = float(qb) / 256.0f;
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    const F4 t00 = ld(x0, y0), t10 = ld(x1, y0), t01 = ld(x0, y1), t11 = ld(x1, y1);
    return ((t00 * w00 + t10 * w10) + t01 * w01) + t11 * w11;
}

// Texel a pixel-centre fetch `texture(img, (vec2(s) + 0.5) / N)` lands on under the pinned sampler, for one axis.
// `exact` is cleared when the quantised weight is not 0 or 1 (then the caller must take the generic bilinear path);
// it never is at real frame sizes (tests/test_post_cpu.py::test_texel_centre_fetches_are_exact).
BLOKB_HD int centre_texel(int s, int N, bool& exact) {
    const float b = ((float(s) + 0.5f) / float(N)) * float(N) - 0.5f;
    const float f = floorf(b);
    const int q = int((b - f) * 256.0f + 0.5f);
    exact = exact && (q == 0 || q == 256);
    const float t = q ? f + 1.0f : f;
    return !(t >= 0.0f) ? 0 : (t > float(N - 1) ? N - 1 : int(t));
}
// normalize() of a G-buffer normal. Voxel faces give exactly unit axis vectors: dot == 1 -> v * (1 / sqrt(1)) == v bit for bit.
BLOKB_HD F3 normalize_gbuf(F3 v) {
    const float d = dot3(v, v);
    return d == 1.0f ? v : v * (1.0f / sqrtf(d));
}

// temporal_reproject.comp main() + variance.comp main() for pixel (x, y).
BLOKB_HD void temporal_variance_pixel(const PostFrame& P, const TemporalIO& io, int x, int y) {
    const int W = (int)P.width, H = (int)P.height;
    const size_t idx = size_t(y) * W + x;
    const F4 cc = ld_rgba32f(io.color, W, x, y);
    const F3 currentColor = f3(cc.x, cc.y, cc.z);
    const F4 wpd = ld_rgba32f(io.worldPos, W, x, y);
    const F4 nrd = ld_rgba16f(io.normalRough, W, x, y);
    const F3 worldPos = f3(wpd.x, wpd.y, wpd.z);
    const float depth = wpd.w;
    const F3 normal = normalize_gbuf(f3(nrd.x, nrd.y, nrd.z));
    const uint32_t mvb = __ldg(reinterpret_cast<const uint32_t*>(io.motion) + idx);
    const float mvx = h2f(mvb & 0xFFFFu), mvy = h2f(mvb >> 16);

    // ---- one walk over the 3x3 neighbourhood feeds both shaders -----------------------------------------------
    // temporal_reproject.comp:118-183 (cross-bilateral YCoCg statistics) and variance.comp:58-99 (spatial variance)
    F3 m1 = f3(0.0f), m2 = f3(0.0f), minVal = f3(1e10f), maxVal = f3(-1e10f);
    float totalWeight = 0.0f;
    float vm1 = 0.0f, vm2 = 0.0f, vTotal = 0.0f;
    for (int dy = -1; dy <= 1; dy++)
        for (int dx = -1; dx <= 1; dx++) {
            const int sx = iclamp(x + dx, 0, W - 1), sy = iclamp(y + dy, 0, H - 1);
            const F4 swp = ld_rgba32f(io.worldPos, W, sx, sy);
            const F4 snd = ld_rgba16f(io.normalRough, W, sx, sy);
            const F4 scd = ld_rgba32f(io.color, W, sx, sy);
            const F3 sampleColor = f3(scd.x, scd.y, scd.z);
            const float depthDiff = fabsf(depth - swp.w);
            {   // temporal statistics: un-normalised sample normal, hard thresholds
                const float normalDot = dot3(normal, f3(snd.x, snd.y, snd.z));
                const float depthWeight = depthDiff < (depth * 0.02f + 0.1f) ? 1.0f : 0.0f;
                const float normalWeight = normalDot > 0.9f ? 1.0f : 0.0f;
                const float weight = depthWeight * normalWeight;
                if (weight > 0.0f) {
                    const F3 s = rgb_to_ycocg(sampleColor);
                    m1 = m1 + s * weight;
                    m2 = m2 + s * s * weight;
                    minVal = gmin3(minVal, s);
                    maxVal = gmax3(maxVal, s);
                    totalWeight += weight;
                }
            }
            {   // spatial variance: normalised sample normal, Gaussian depth weight
                const float normalDot = dot3(normal, normalize_gbuf(f3(snd.x, snd.y, snd.z)));
                const float depthWeight = expf(-depthDiff * depthDiff / (0.5f * 0.5f));
                const float normalWeight = normalDot > 0.9f ? 1.0f : 0.0f;
                const float weight = depthWeight * normalWeight;
                if (weight > 0.01f) {
                    const float l = lum709(sampleColor);
                    vm1 += l * weight;
                    vm2 += l * l * weight;
                    vTotal += weight;
                }
            }
        }

    // ---- temporal_reproject.comp:205-316 ------------------------------------------------------------------------
    const float cuX = (float(x) + 0.5f) / float(P.width), cuY = (float(y) + 0.5f) / float(P.height);
    float puX, puY;
    if (sqrtf(mvx * mvx + mvy * mvy) > 0.0001f) {
        puX = cuX - mvx;
        puY = cuY - mvy;
    } else {
        const F4 pc = mat_mul(P.prevViewProj, worldPos.x, worldPos.y, worldPos.z, 1.0f);
        puX = (pc.x / pc.w) * 0.5f + 0.5f;
        puY = (pc.y / pc.w) * 0.5f + 0.5f;
    }
    F3 outputColor = currentColor;
    const float lum = lum709(currentColor);
    float outM0 = lum, outM1 = lum * lum;
    float outputHistoryLength = 1.0f;
    const bool validReprojection = (puX >= 0.0f && puX <= 1.0f && puY >= 0.0f && puY <= 1.0f) && P.frameCount > 0u;
    if (validReprojection) {
        const F4 hc = tex_linear<false>(io.prevHistory, W, H, puX, puY);
        const F3 historyColor = f3(hc.x, hc.y, hc.z);
        const int pcx = iclamp(int(puX * float(P.width)), 0, W - 1), pcy = iclamp(int(puY * float(P.height)), 0, H - 1);
        const F4 pwd = ld_rgba32f(io.prevWorldPos, W, pcx, pcy);
        const F4 pnd = ld_rgba16f(io.prevNormal, W, pcx, pcy);
        const float prevDepth = pwd.w;
        const F3 prevNormal = normalize_gbuf(f3(pnd.x, pnd.y, pnd.z));
        const float absoluteDepthThreshold = P.depthThreshold * depth + 0.5f;
        const bool depthValid = fabsf(depth - prevDepth) < absoluteDepthThreshold;
        const bool normalValid = dot3(normal, prevNormal) > P.normalThreshold;
        const bool worldPosValid = length3(worldPos - f3(pwd.x, pwd.y, pwd.z)) < 2.0f;
        if (depthValid && normalValid && worldPosValid) {
            const size_t pidx = size_t(pcy) * W + pcx;
            const float2 pm = __ldg(reinterpret_cast<const float2*>(io.prevMoments) + pidx);
            const float prevHistLen = h2f(__ldg(io.prevLength + pidx));
            // statistics -> clip box (:161-182)
            F3 mean, stdDev;
            if (totalWeight > 0.0f) {
                mean = m1 / totalWeight;
                stdDev = sqrt3(gmax3(m2 / totalWeight - mean * mean, f3(0.0f)));
            } else {
                mean = rgb_to_ycocg(currentColor);
                stdDev = f3(0.1f);
                minVal = mean;
                maxVal = mean;
            }
            const float gamma = P.varianceClipGamma;
            F3 minC = mean - gamma * stdDev, maxC = mean + gamma * stdDev;
            minC = gmax3(minC, minVal - f3(0.05f));
            maxC = gmin3(maxC, maxVal + f3(0.05f));
            // clipToAABB (:93-106)
            const F3 hy = rgb_to_ycocg(historyColor);
            const F3 center = 0.5f * (minC + maxC), extents = 0.5f * (maxC - minC);
            const F3 offset = hy - center;
            const F3 unit = offset / gmax3(extents, f3(0.0001f));
            const float maxComponent = gmax(gmax(fabsf(unit.x), fabsf(unit.y)), fabsf(unit.z));
            const F3 clippedY = maxComponent > 1.0f ? center + offset / maxComponent : hy;
            const F3 clippedHistory = gmax3(ycocg_to_rgb(clippedY), f3(0.0f));
            // blend (:271-310)
            const float historyLen = prevHistLen + 1.0f;
            float alpha = P.temporalAlpha;
            const float historyFactor = 1.0f / gmax(historyLen, 1.0f);
            alpha = gmax(alpha, historyFactor);
            const float lumCurrent = lum709(currentColor), lumHistory = lum709(clippedHistory);
            const float lumDiff = fabsf(lumCurrent - lumHistory) / gmax(lumCurrent + lumHistory + 0.01f, 0.01f);
            alpha = gmix(alpha, gmin(alpha + 0.2f, 0.5f), lumDiff * 0.3f);
            alpha = gclamp(alpha, P.temporalAlpha, 1.0f);
            outputColor = mix3(clippedHistory, currentColor, alpha);
            const float newLum = lum709(currentColor);
            const float momentAlpha = gmax(P.momentAlpha, historyFactor);
            outM0 = gmix(pm.x, newLum, momentAlpha);
            outM1 = gmix(pm.y, newLum * newLum, momentAlpha);
            outputHistoryLength = gmin(historyLen, 64.0f);
        }
    }
    outputColor = f3(gclamp(outputColor.x, 0.0f, 100.0f), gclamp(outputColor.y, 0.0f, 100.0f), gclamp(outputColor.z, 0.0f, 100.0f));
    outM0 = gclamp(outM0, 0.0f, 10000.0f);
    outM1 = gclamp(outM1, 0.0f, 10000.0f);
    reinterpret_cast<float4*>(io.outHistory)[idx] = make_float4(outputColor.x, outputColor.y, outputColor.z, 1.0f);
    reinterpret_cast<float2*>(io.outMoments)[idx] = make_float2(outM0, outM1);
    const uint32_t lenBits = f2h(outputHistoryLength);
    io.outLength[idx] = (uint16_t)lenBits;

    // ---- variance.comp:109-144 (reads back what was just stored) --------------------------------------------------
    const float historyLength = h2f(lenBits);
    const float temporalVariance = gmax(outM1 - outM0 * outM0, 0.0f);
    float spatialVariance = 0.0f;
    if (vTotal > 0.0f) {
        const float mean = vm1 / vTotal;
        spatialVariance = gmax(vm2 / vTotal - mean * mean, 0.0f);
    }
    const float minHistLen = float(P.minHistoryLength < 4 ? 4 : P.minHistoryLength);
    float historyWeight = gclamp((historyLength - 1.0f) / minHistLen, 0.0f, 1.0f);
    historyWeight = historyWeight * historyWeight;
    float variance = gmix(spatialVariance, temporalVariance, historyWeight);
    if (historyLength < minHistLen) variance *= gmix(P.varianceBoost, 1.0f, historyLength / minHistLen);
    io.outVariance[idx] = gmax(variance, 0.0001f);
}

// atrous.comp main() for pixel (x, y): in -> out with step `stepSize`.
// Same results as the shader for every input; the work is ordered so that the common voxel case is cheap: the 5 + 5
// distinct tap columns / rows are resolved to texels once (instead of 25 sampler evaluations), a tap is dropped as soon
// as one factor of its weight is known to be zero (weight < 0.001 discards it in the shader too), and unit normals
// skip normalisation and the normal-weight division ((1 - 0.9) / (1 - 0.9) == 1).
BLOKB_HD void atrous_pixel(const PostFrame& P, int stepSize, const float* inColor, const float* variance, const float* worldPosImg,
                           const uint16_t* normalRough, float* outColor, int x, int y) {
    const int W = (int)P.width, H = (int)P.height;
    const size_t idx = size_t(y) * W + x;
    const F4 c4 = tex_linear<false>(inColor, W, H, (float(x) + 0.5f) / float(P.width), (float(y) + 0.5f) / float(P.height));
    const F3 centerColor = f3(c4.x, c4.y, c4.z);
    const F4 cw = ld_rgba32f(worldPosImg, W, x, y);
    const float centerDepth = cw.w;
    if (centerDepth > 9000.0f) {  // sky
        reinterpret_cast<float4*>(outColor)[idx] = make_float4(centerColor.x, centerColor.y, centerColor.z, 1.0f);
        return;
    }
    const F4 cn4 = ld_rgba16f(normalRough, W, x, y);
    const float centerVariance = __ldg(variance + idx);
    const F3 centerWorldPos = f3(cw.x, cw.y, cw.z);
    const F3 centerNormal = normalize_gbuf(f3(cn4.x, cn4.y, cn4.z));
    // computeColorWeight's sigma depends on the centre only (:85-94)
    const float csigma = 0.01f + P.phiColor * sqrtf(gmax(centerVariance, 0.0f));
    const float cden = 2.0f * csigma * csigma + 1e-6f;
    const float dsigma = P.phiDepth * float(stepSize) + 0.1f;  // (:131)
    const float dden = dsigma * dsigma + 1e-6f;
    const float dcut = dsigma * 2.0f;
    auto kern_w = [](int o) { return o == 0 ? 1.0f : ((o == 1 || o == -1) ? 2.0f / 3.0f : 1.0f / 6.0f); };  // kernel[abs(o)] (:58)
    int sxs[5], txs[5];
    bool exact = true;
#pragma unroll
    for (int k = 0; k < 5; k++) {
        sxs[k] = iclamp(x + (k - 2) * stepSize, 0, W - 1);
        txs[k] = centre_texel(sxs[k], W, exact);
        centre_texel(iclamp(y + (k - 2) * stepSize, 0, H - 1), H, exact);
    }
    F3 sumColor = f3(0.0f);
    float sumWeight = 0.0f;
    auto tap = [&](int ox, int oy, int sx, int sy, int txl, int tyl, bool fast) {
        const F4 sw = ld_rgba32f(worldPosImg, W, sx, sy);
        if (sw.w > 9000.0f) return;
        const uint2 nb = __ldg(reinterpret_cast<const uint2*>(normalRough) + (unsigned(sy) * unsigned(W) + unsigned(sx)));
        const F3 sampleNormal = normalize_gbuf(f3(h2f(nb.x & 0xFFFFu), h2f(nb.x >> 16), h2f(nb.y & 0xFFFFu)));
        float normalWeight;
        {
            const float dp = gmax(dot3(centerNormal, sampleNormal), 0.0f);
            if (dp < 0.9f) normalWeight = 0.0f;
            else if (dp == 1.0f) normalWeight = 1.0f;
            else {
                const float t = (dp - 0.9f) / (1.0f - 0.9f);
                normalWeight = t * t;
            }
        }
        const float depthDiff = fabsf(centerDepth - sw.w);
        const float planeDistance = fabsf(dot3(f3(sw.x, sw.y, sw.z) - centerWorldPos, centerNormal));
        const float eff = gmax(depthDiff * 0.1f, planeDistance);
        const bool depthCut = eff > dcut;
        const F4 s4 = fast ? ld_rgba32f(inColor, W, txl, tyl)
                           : tex_linear<false>(inColor, W, H, (float(sx) + 0.5f) / float(P.width), (float(sy) + 0.5f) / float(P.height));
        const F3 sampleColor = f3(s4.x, s4.y, s4.z);
        const F3 diff = centerColor - sampleColor;
        const float cq = dot3(diff, diff) / cden;
        // a zero factor with finite partners makes the product (+)0 < 0.001: the shader `continue`s. (cq is NaN only for a NaN colour.)
        if ((normalWeight == 0.0f || depthCut) && cq == cq) return;
        const float colorWeight = 1.0f / (1.0f + cq);
        const float depthWeight = depthCut ? 0.0f : 1.0f / (1.0f + eff * eff / dden);
        const float kernelWeight = kern_w(ox) * kern_w(oy);
        const float weight = kernelWeight * colorWeight * normalWeight * depthWeight;
        if (weight < 0.001f) return;
        sumColor = sumColor + sampleColor * weight;
        sumWeight += weight;
    };
    if (exact) {
#pragma unroll 1
        for (int oy = -2; oy <= 2; oy++) {  // rows stay a loop (instruction-cache footprint), the five columns are unrolled
            const int sy = iclamp(y + oy * stepSize, 0, H - 1);
            bool e2 = true;
            const int ty = centre_texel(sy, H, e2);
#pragma unroll
            for (int ox = -2; ox <= 2; ox++) tap(ox, oy, sxs[ox + 2], sy, txs[ox + 2], ty, true);
        }
    } else {
#pragma unroll 1
        for (int i = 0; i < 25; i++) {
            const int ox = i % 5 - 2, oy = i / 5 - 2;
            tap(ox, oy, iclamp(x + ox * stepSize, 0, W - 1), iclamp(y + oy * stepSize, 0, H - 1), 0, 0, false);
        }
    }
    F3 o = sumWeight > 0.01f ? sumColor / sumWeight : centerColor;
    o = gmax3(o, f3(0.0f));
    reinterpret_cast<float4*>(outColor)[idx] = make_float4(o.x, o.y, o.z, 1.0f);
}

// tonemap.comp main() body on an HDR texel; returns the rgba8 word it stores.
BLOKB_HD uint32_t tonemap_texel(const PostFrame& P, F3 hdr) {
    hdr = hdr * P.exposure;
    const F3 hdrOriginal = hdr;
    F3 ldr;
    if (P.tonemapOperator == 0) {  // neutralTonemap (:96-107)
        const float peak = gmax(gmax(hdr.x, hdr.y), hdr.z);
        if (peak <= 1.0f) ldr = hdr;
        else {
            const float compressed = 1.0f - expf(-(peak - 1.0f));
            ldr = hdr * ((1.0f + compressed) / peak);
        }
    } else {  // khronosPbrNeutral (:74-93)
        const float startCompression = 0.8f - 0.04f, desaturation = 0.15f;
        const float mn = gmin(hdr.x, gmin(hdr.y, hdr.z));
        const float offset = mn < 0.08f ? mn - 6.25f * mn * mn : 0.04f;
        hdr = hdr - f3(offset);
        const float peak = gmax(hdr.x, gmax(hdr.y, hdr.z));
        if (peak < startCompression) ldr = hdr;
        else {
            const float d = 1.0f - startCompression;
            const float newPeak = 1.0f - d * d / (peak + d - startCompression);
            hdr = hdr * (newPeak / peak);
            const float g = 1.0f - 1.0f / (desaturation * (peak - newPeak) + 1.0f);
            ldr = mix3(hdr, f3(newPeak), g);
        }
    }
    if (P.saturationBoost > 1.0f) {  // postTonemapSaturationBoost (:51-70)
        const float hdrLuma = lum709(hdrOriginal);
        const float hdrSat = hdrLuma > 0.0001f ? length3(hdrOriginal - f3(hdrLuma)) / hdrLuma : 0.0f;
        const float ldrLuma = lum709(ldr);
        const float ldrSat = ldrLuma > 0.0001f ? length3(ldr - f3(ldrLuma)) / ldrLuma : 0.0f;
        if (ldrSat > 0.0001f && ldrLuma > 0.01f) {
            const float satRatio = gmin(hdrSat / gmax(ldrSat, 0.001f), 2.0f);
            const float recovery = gmix(1.0f, satRatio, P.saturationBoost - 1.0f);
            ldr = mix3(f3(ldrLuma), ldr, gmin(recovery, 1.5f));
        }
    } else if (P.saturationBoost < 1.0f && P.saturationBoost > 0.0f) {
        ldr = mix3(f3(lum709(ldr)), ldr, P.saturationBoost);
    }
    return pack_unorm8(ldr.x) | pack_unorm8(ldr.y) << 8 | pack_unorm8(ldr.z) << 16 | 255u << 24;
}

// taa.comp main() for pixel (x, y); with TONEMAP the tonemap pass is applied to the value TAA would have stored
// in taaOutput (tonemap.comp samples it at the texel centre) and taaOutput itself is not materialised.
template <bool TAA, bool TONEMAP>
BLOKB_HD void taa_tonemap_pixel(const PostFrame& P, const float* currentColor, const float* prevHistory, const uint16_t* motionImg,
                                float* outHistory, float* outColor, uint32_t* outLdr, int x, int y) {
    const int W = (int)P.width, H = (int)P.height;
    const size_t idx = size_t(y) * W + x;
    const F4 cs = ld_rgba32f(currentColor, W, x, y);
    F3 outputResult = f3(cs.x, cs.y, cs.z);
    if constexpr (TAA) {
        const F3 current = outputResult;
        const float uvx = (float(x) + 0.5f) / float(W), uvy = (float(y) + 0.5f) / float(H);
        const uint32_t mvb = __ldg(reinterpret_cast<const uint32_t*>(motionImg) + idx);
        const float mx = h2f(mvb & 0xFFFFu), my = h2f(mvb >> 16);
        const float pux = uvx - mx, puy = uvy - my;
        const bool validHistory = pux >= 0.0f && pux <= 1.0f && puy >= 0.0f && puy <= 1.0f;
        const F4 h4 = tex_linear<false>(prevHistory, W, H, pux, puy);
        const F3 history = f3(h4.x, h4.y, h4.z);
        F3 nMin = f3(1e10f), nMax = f3(-1e10f), nSum = f3(0.0f), nSumSq = f3(0.0f);
        for (int dy = -1; dy <= 1; dy++)
            for (int dx = -1; dx <= 1; dx++) {
                const F4 s4 = ld_rgba32f(currentColor, W, iclamp(x + dx, 0, W - 1), iclamp(y + dy, 0, H - 1));
                const F3 s = rgb_to_ycocg(f3(s4.x, s4.y, s4.z));
                nMin = gmin3(nMin, s);
                nMax = gmax3(nMax, s);
                nSum = nSum + s;
                nSumSq = nSumSq + s * s;
            }
        const F3 mean = nSum / 9.0f;
        const F3 variance = (nSumSq / 9.0f) - (mean * mean);
        const F3 stdDev = sqrt3(gmax3(variance, f3(0.0f)));
        const F3 historyY = rgb_to_ycocg(history);
        // varianceClip (:92-103) + clipToAABB (:80-89)
        const float gamma = 1.5f;
        const F3 minC = gmax3(mean - gamma * stdDev, nMin), maxC = gmin3(mean + gamma * stdDev, nMax);
        const F3 center = 0.5f * (maxC + minC), extents = 0.5f * (maxC - minC);
        const F3 offset = historyY - center;
        const F3 ts = abs3(extents) / gmax3(abs3(offset), f3(0.0001f));
        const float t = gclamp(gmin(gmin(ts.x, ts.y), ts.z), 0.0f, 1.0f);
        const F3 clippedY = center + offset * t;
        const F3 clippedHistory = ycocg_to_rgb(clippedY);
        const float vx = mx * float(W), vy = my * float(H);
        const float velocityFactor = gclamp(sqrtf(vx * vx + vy * vy) / 10.0f, 0.0f, 1.0f);
        float feedback = gmix(P.feedbackMax, P.feedbackMin, velocityFactor);
        if (!validHistory || P.frameCount == 0u) feedback = 0.0f;
        const float clipDist = length3(clippedY - historyY);
        feedback *= 1.0f - gclamp(clipDist * 2.0f, 0.0f, 0.5f);
        const F3 result = mix3(current, clippedHistory, feedback);
        const F3 sharpened = current + 0.1f * (current - ycocg_to_rgb(mean));
        outputResult = mix3(sharpened, clippedHistory, feedback);
        reinterpret_cast<float4*>(outHistory)[idx] = make_float4(result.x, result.y, result.z, 1.0f);
    }
    if constexpr (TONEMAP) outLdr[idx] = tonemap_texel(P, outputResult);
    else reinterpret_cast<float4*>(outColor)[idx] = make_float4(outputResult.x, outputResult.y, outputResult.z, cs.w);
}

// One axis of the pinned sampler for an arbitrary coordinate (sharpen.comp's uv +- texelSize taps).
BLOKB_HD int axis_texel(float u, int N, bool& exact) {
    const float b = u * float(N) - 0.5f;
    const float f = floorf(b);
    const int q = int((b - f) * 256.0f + 0.5f);
    exact = exact && (q == 0 || q == 256);
    const float t = q ? f + 1.0f : f;
    return !(t >= 0.0f) ? 0 : (t > float(N - 1) ? N - 1 : int(t));
}

// sharpen.comp main() for pixel (x, y). lut255[i] = float(i) / 255.0f (the unorm8 decode, computed once per block with
// the same IEEE division); the 3 + 3 tap columns / rows are resolved to texels once.
BLOKB_HD void sharpen_pixel(const PostFrame& P, const float* lut255, const uint8_t* in, uint32_t* out, int x, int y) {
    const int W = (int)P.width, H = (int)P.height;
    const float tx = 1.0f / float(W), ty = 1.0f / float(H);
    const float u = (float(x) + 0.5f) / float(W), v = (float(y) + 0.5f) / float(H);
    bool exact = true;
    int cx[3], cy[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        cx[k] = axis_texel(u + tx * float(k - 1), W, exact);
        cy[k] = axis_texel(v + ty * float(k - 1), H, exact);
    }
    auto tap = [&](int ox, int oy) {
        if (exact) {
            const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(in) + (size_t(cy[oy + 1]) * W + cx[ox + 1]));
            return f3(lut255[w & 255u], lut255[(w >> 8) & 255u], lut255[(w >> 16) & 255u]);
        }
        const F4 t = tex_linear<true>(in, W, H, u + tx * float(ox), v + ty * float(oy));
        return f3(t.x, t.y, t.z);
    };
    const F3 a = tap(-1, -1), b = tap(0, -1), c = tap(1, -1);
    const F3 d = tap(-1, 0), e = tap(0, 0), f = tap(1, 0);
    const F3 g = tap(-1, 1), h = tap(0, 1), i = tap(1, 1);
    const F3 blur = (1.0f * (a + c + g + i) + 2.0f * (b + d + f + h) + 4.0f * e) / 16.0f;
    const F3 detail = e - blur;
    const float intensity = P.sharpenStrength * 3.0f;
    const F3 r = e + detail * intensity;
    out[size_t(y) * W + x] = pack_unorm8(r.x) | pack_unorm8(r.y) << 8 | pack_unorm8(r.z) << 16 | 255u << 24;
}

}  // namespace post
}  // namespace blokb
