// Host-visible side of the denoiser / post-process kernels (post.cu): parameter blocks and launchers.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/blokb200.h"

namespace blokb {
namespace post {

// FrameUBO fields the compute shaders read (resources.hpp:103-150) + push constants of the three post passes.
struct PostFrame {
    float prevViewProj[16];
    uint32_t frameCount, width, height;
    float temporalAlpha, momentAlpha, varianceClipGamma, depthThreshold, normalThreshold, varianceBoost;
    int32_t minHistoryLength;
    float phiColor, phiNormal, phiDepth;                 // AtrousPC (renderer_denoising.cpp:812-817)
    float feedbackMin, feedbackMax;                      // TAAPushConstants
    float exposure, saturationBoost, whitePoint;         // TonemapPushConstants
    int32_t tonemapOperator;
    float sharpenStrength;                               // SharpenPushConstants
};

struct TemporalIO {
    const float* color;          // gbuffer.color            rgba32f
    const float* worldPos;       // gbuffer.worldPosition    rgba32f (xyz, hitT)
    const uint16_t* normalRough; // gbuffer.normalRoughness  rgba16f
    const uint16_t* motion;      // gbuffer.motionVectors    rg16f
    const float* prevHistory;    // previousHistory()        rgba32f, sampled
    const float* prevMoments;    // previousMoments()        rg32f
    const uint16_t* prevLength;  // previousHistoryLength()  r16f
    const float* prevWorldPos;   // previousWorldPosition()
    const uint16_t* prevNormal;  // previousNormalRoughness()
    float* outHistory;           // currentHistory()
    float* outMoments;           // currentMoments()
    uint16_t* outLength;         // currentHistoryLength()
    float* outVariance;          // gbuffer.variance          r32f
};


// PostFrame from the FrameUBO (what fillFrameUBO put there) and the post settings (what the reference pushes as constants).
inline PostFrame make_post_frame(const blokb_frame_ubo& u, const blokb_post_settings& s) {
    PostFrame P{};
    for (int i = 0; i < 16; i++) P.prevViewProj[i] = u.prevViewProj[i];
    P.frameCount = u.frameCount; P.width = u.screenWidth; P.height = u.screenHeight;
    P.temporalAlpha = u.temporalAlpha; P.momentAlpha = u.momentAlpha; P.varianceClipGamma = u.varianceClipGamma;
    P.depthThreshold = u.depthThreshold; P.normalThreshold = u.normalThreshold; P.varianceBoost = u.varianceBoost;
    P.minHistoryLength = u.minHistoryLength;
    P.phiColor = u.phiColor; P.phiNormal = u.phiNormal; P.phiDepth = u.phiDepth;
    P.feedbackMin = s.feedbackMin; P.feedbackMax = s.feedbackMax;
    P.exposure = s.exposure; P.saturationBoost = s.saturationBoost; P.whitePoint = s.whitePoint;
    P.tonemapOperator = s.tonemapOperator; P.sharpenStrength = s.sharpenStrength;
    return P;
}

cudaError_t launch_temporal_variance(const PostFrame& P, const TemporalIO& io, cudaStream_t st);
cudaError_t launch_atrous(const PostFrame& P, int stepSize, const float* in, const float* variance, const float* worldPos,
                          const uint16_t* normalRough, float* out, cudaStream_t st);
cudaError_t launch_taa_tonemap(const PostFrame& P, bool taa, bool tonemap, const float* current, const float* prevHistory,
                               const uint16_t* motion, float* outHistory, float* outColor, uint32_t* outLdr, cudaStream_t st);
cudaError_t launch_sharpen(const PostFrame& P, const uint8_t* in, uint32_t* out, cudaStream_t st);
}  // namespace post
}  // namespace blokb
