// Denoiser + post-process kernels for sm_100a. One thread per pixel, 32x8-pixel blocks
// (a warp covers a 32x1 row segment: every 16-B / 8-B / 4-B image access of a warp is one contiguous run).
// The per-pixel arithmetic lives in post.cuh; this file is launch geometry only.
#include "post.cuh"

namespace blokb {
namespace post {
namespace {
constexpr int kBX = 32, kBY = 8;
static_assert(kBX * kBY == 256, "k_sharpen fills its 256-entry unorm8 table with one entry per thread");
inline dim3 grid_for(const PostFrame& P) { return dim3((P.width + kBX - 1) / kBX, (P.height + kBY - 1) / kBY); }

__global__ void __launch_bounds__(kBX* kBY, 4) k_temporal_variance(const __grid_constant__ PostFrame P, const __grid_constant__ TemporalIO io) {
    const int x = blockIdx.x * kBX + threadIdx.x, y = blockIdx.y * kBY + threadIdx.y;
    if (x >= (int)P.width || y >= (int)P.height) return;
    temporal_variance_pixel(P, io, x, y);
}
#ifndef BLOKB_ATROUS_MIN_BLOCKS
#define BLOKB_ATROUS_MIN_BLOCKS 5  // measured at 4K: 3 blocks/SM 4.43 ms, 4 (63 regs) 3.77, none (58 regs) 3.83, 5 (48 regs) 3.49, 6 (40 regs) 3.48 per denoise
#endif
#if BLOKB_ATROUS_MIN_BLOCKS > 0
#define BLOKB_ATROUS_BOUNDS __launch_bounds__(kBX* kBY, BLOKB_ATROUS_MIN_BLOCKS)
#else
#define BLOKB_ATROUS_BOUNDS __launch_bounds__(kBX* kBY)
#endif
__global__ void BLOKB_ATROUS_BOUNDS k_atrous(const __grid_constant__ PostFrame P, int stepSize, const float* __restrict__ in,
                                                     const float* __restrict__ variance, const float* __restrict__ worldPos,
                                                     const uint16_t* __restrict__ normalRough, float* __restrict__ out) {
    const int x = blockIdx.x * kBX + threadIdx.x, y = blockIdx.y * kBY + threadIdx.y;
    if (x >= (int)P.width || y >= (int)P.height) return;
    atrous_pixel(P, stepSize, in, variance, worldPos, normalRough, out, x, y);
}
template <bool TAA, bool TONEMAP>
__global__ void __launch_bounds__(kBX* kBY) k_taa_tonemap(const __grid_constant__ PostFrame P, const float* __restrict__ current,
                                                          const float* __restrict__ prevHistory, const uint16_t* __restrict__ motion,
                                                          float* __restrict__ outHistory, float* __restrict__ outColor, uint32_t* __restrict__ outLdr) {
    const int x = blockIdx.x * kBX + threadIdx.x, y = blockIdx.y * kBY + threadIdx.y;
    if (x >= (int)P.width || y >= (int)P.height) return;
    taa_tonemap_pixel<TAA, TONEMAP>(P, current, prevHistory, motion, outHistory, outColor, outLdr, x, y);
}
__global__ void __launch_bounds__(kBX* kBY) k_sharpen(const __grid_constant__ PostFrame P, const uint8_t* __restrict__ in, uint32_t* __restrict__ out) {
    __shared__ float lut255[256];
    lut255[threadIdx.y * kBX + threadIdx.x] = float(threadIdx.y * kBX + threadIdx.x) / 255.0f;
    __syncthreads();
    const int x = blockIdx.x * kBX + threadIdx.x, y = blockIdx.y * kBY + threadIdx.y;
    if (x >= (int)P.width || y >= (int)P.height) return;
    sharpen_pixel(P, lut255, in, out, x, y);
}
}  // namespace

cudaError_t launch_temporal_variance(const PostFrame& P, const TemporalIO& io, cudaStream_t st) {
    k_temporal_variance<<<grid_for(P), dim3(kBX, kBY), 0, st>>>(P, io);
    return cudaGetLastError();
}
cudaError_t launch_atrous(const PostFrame& P, int stepSize, const float* in, const float* variance, const float* worldPos,
                          const uint16_t* normalRough, float* out, cudaStream_t st) {
    k_atrous<<<grid_for(P), dim3(kBX, kBY), 0, st>>>(P, stepSize, in, variance, worldPos, normalRough, out);
    return cudaGetLastError();
}
cudaError_t launch_taa_tonemap(const PostFrame& P, bool taa, bool tonemap, const float* current, const float* prevHistory,
                               const uint16_t* motion, float* outHistory, float* outColor, uint32_t* outLdr, cudaStream_t st) {
    const dim3 g = grid_for(P), b(kBX, kBY);
    if (taa && tonemap) k_taa_tonemap<true, true><<<g, b, 0, st>>>(P, current, prevHistory, motion, outHistory, outColor, outLdr);
    else if (taa) k_taa_tonemap<true, false><<<g, b, 0, st>>>(P, current, prevHistory, motion, outHistory, outColor, outLdr);
    else if (tonemap) k_taa_tonemap<false, true><<<g, b, 0, st>>>(P, current, prevHistory, motion, outHistory, outColor, outLdr);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}
cudaError_t launch_sharpen(const PostFrame& P, const uint8_t* in, uint32_t* out, cudaStream_t st) {
    k_sharpen<<<grid_for(P), dim3(kBX, kBY), 0, st>>>(P, in, out);
    return cudaGetLastError();
}
}  // namespace post
}  // namespace blokb
