// TEST INFRASTRUCTURE ONLY — part of oracle/_ref/libblokref.so
//
// CPU harness around the reference's own *compute* shaders, compiled from where they
// lie under $BLOK_REF after the build-time text rewrite of oracle/glsl2cpp.py (kind "comp"):
//   assets/shaders/temporal_reproject.comp, variance.comp, atrous.comp   (Denoiser, renderer_denoising.cpp:714-769)
//   assets/shaders/taa.comp, tonemap.comp, sharpen.comp                  (PostProcess, renderer_postprocess.cpp:505-558)
// This file supplies only what the Vulkan driver supplies: gl_GlobalInvocationID, formatted image load/store,
// the combined-image-sampler fetch and the dispatch loop. Which image goes to which binding and in which order the
// passes run is decided by the caller (oracle/pyoracle.py: RefPostChain, following updateDescriptorSets /
// denoise / process of the reference).
//
// Driver behaviour that is not in the reference tree and is pinned here ("parity unpinned"):
//   * float -> half on imageStore to r16f/rg16f/rgba16f: round-to-nearest-even;
//   * unorm8 store: round(clamp(v,0,1) * 255) (Vulkan's conversion rule); load: u8 / 255.0f;
//   * texture() with the reference's linear / clamp-to-edge sampler: unnormalised coordinate u*W - 0.5 in fp32,
//     i = floor, fractional weight quantised to 8 bits round-to-nearest (the 1.8 fixed-point weights NVIDIA
//     documents for its texture units), result ((T00*w00 + T10*w10) + T01*w01) + T11*w11 in fp32.
#include <glm.hpp>

#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace compenv {
using namespace glm;
typedef uint32_t uint;

enum Fmt { RGBA32F = 0, RG32F = 1, R32F = 2, RGBA16F = 3, RG16F = 4, R16F = 5, RGBA8 = 6 };
struct image2D {
    void* data = nullptr;
    int width = 0, height = 0, fmt = 0;
};
struct sampler2D {
    image2D img;
};
static thread_local uvec3 gl_GlobalInvocationID;

static inline float half_to_float(uint16_t h) {
    uint32_t s = (h >> 15) & 1u, e = (h >> 10) & 31u, m = h & 1023u, u;
    if (e == 0) {
        if (m == 0) u = s << 31;
        else {
            int sh = 0;
            while (!(m & 1024u)) { m <<= 1; sh++; }
            u = (s << 31) | uint32_t(127 - 15 - sh + 1) << 23 | ((m & 1023u) << 13);
        }
    } else if (e == 31) u = (s << 31) | 0x7F800000u | (m << 13);
    else u = (s << 31) | ((e + 112u) << 23) | (m << 13);
    float f; std::memcpy(&f, &u, 4); return f;
}
static inline uint16_t float_to_half(float f) {  // round-to-nearest-even, IEEE binary16
    uint32_t u; std::memcpy(&u, &f, 4);
    uint32_t s = (u >> 16) & 0x8000u, a = u & 0x7FFFFFFFu;
    if (a >= 0x7F800000u) return uint16_t(s | 0x7C00u | (a > 0x7F800000u ? 0x200u | ((a >> 13) & 1023u) : 0u));
    if (a >= 0x477FF000u) return uint16_t(s | 0x7C00u);                  // rounds to >= 65520 -> inf
    if (a < 0x33000001u) return uint16_t(s);                             // <= 2^-25 -> 0
    int e = int(a >> 23) - 127;
    uint32_t m = (a & 0x7FFFFFu) | 0x800000u;
    int shift = e < -14 ? 13 + (-14 - e) : 13;
    uint32_t q = m >> shift, rem = m & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (q & 1u))) q++;
    uint32_t h = e < -14 ? q : (uint32_t(e + 15) << 10) + (q - 1024u);   // mantissa carry propagates into the exponent
    return uint16_t(s | h);
}
static inline int nchan(int fmt) { return fmt == RGBA32F || fmt == RGBA16F || fmt == RGBA8 ? 4 : (fmt == RG32F || fmt == RG16F ? 2 : 1); }

static inline vec4 imageLoad(const image2D& im, ivec2 p) {
    const size_t i = size_t(p.y) * im.width + p.x;
    const int n = nchan(im.fmt);
    float c[4] = {0.f, 0.f, 0.f, 1.f};
    switch (im.fmt) {
        case RGBA32F: case RG32F: case R32F:
            for (int k = 0; k < n; k++) c[k] = ((const float*)im.data)[i * n + k];
            break;
        case RGBA16F: case RG16F: case R16F:
            for (int k = 0; k < n; k++) c[k] = half_to_float(((const uint16_t*)im.data)[i * n + k]);
            break;
        default:
            for (int k = 0; k < 4; k++) c[k] = float(((const uint8_t*)im.data)[i * 4 + k]) / 255.0f;
    }
    return vec4(c[0], c[1], c[2], c[3]);
}
static inline void imageStore(image2D& im, ivec2 p, vec4 v) {
    if (!im.data) return;
    const size_t i = size_t(p.y) * im.width + p.x;
    const int n = nchan(im.fmt);
    const float c[4] = {v.x, v.y, v.z, v.w};
    switch (im.fmt) {
        case RGBA32F: case RG32F: case R32F:
            for (int k = 0; k < n; k++) ((float*)im.data)[i * n + k] = c[k];
            break;
        case RGBA16F: case RG16F: case R16F:
            for (int k = 0; k < n; k++) ((uint16_t*)im.data)[i * n + k] = float_to_half(c[k]);
            break;
        default:
            for (int k = 0; k < 4; k++) {
                float x = c[k] < 0.f ? 0.f : (c[k] > 1.f ? 1.f : c[k]);
                if (!(x == x)) x = 0.f;
                ((uint8_t*)im.data)[i * 4 + k] = (uint8_t)std::floor(x * 255.0f + 0.5f);
            }
    }
}
static inline ivec2 imageSize(const image2D& im) { return ivec2(im.width, im.height); }

static inline vec4 texture(const sampler2D& s, vec2 uv) {
    const image2D& im = s.img;
    const float xb = uv.x * float(im.width) - 0.5f, yb = uv.y * float(im.height) - 0.5f;
    const float fx = std::floor(xb), fy = std::floor(yb);
    int qa = int((xb - fx) * 256.0f + 0.5f), qb = int((yb - fy) * 256.0f + 0.5f);
    // clamp the integer coordinates in float first so that absurd uv (garbage history) cannot overflow int
    auto clampi = [](float f, int hi) { return !(f >= 0.f) ? 0 : (f > float(hi) ? hi : int(f)); };
    const int x0 = clampi(fx, im.width - 1), x1 = clampi(fx + 1.0f, im.width - 1);
    const int y0 = clampi(fy, im.height - 1), y1 = clampi(fy + 1.0f, im.height - 1);
    if (!(xb == xb)) qa = 0;
    if (!(yb == yb)) qb = 0;
    const float a = float(qa) / 256.0f, b = float(qb) / 256.0f;
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    if (qa == 0 && qb == 0) return imageLoad(im, ivec2(x0, y0));
    if (qa == 256 && qb == 0) return imageLoad(im, ivec2(x1, y0));
    if (qa == 0 && qb == 256) return imageLoad(im, ivec2(x0, y1));
    if (qa == 256 && qb == 256) return imageLoad(im, ivec2(x1, y1));
    const vec4 t00 = imageLoad(im, ivec2(x0, y0)), t10 = imageLoad(im, ivec2(x1, y0));
    const vec4 t01 = imageLoad(im, ivec2(x0, y1)), t11 = imageLoad(im, ivec2(x1, y1));
    return ((t00 * w00 + t10 * w10) + t01 * w01) + t11 * w11;
}

namespace temporal {
#include "gen/temporal_reproject.comp.inc"
}
namespace variance {
#include "gen/variance.comp.inc"
}
namespace atrous {
#include "gen/atrous.comp.inc"
}
namespace taa {
#include "gen/taa.comp.inc"
}
namespace tonemap {
#include "gen/tonemap.comp.inc"
}
namespace sharpen {
#include "gen/sharpen.comp.inc"
}
static_assert(sizeof(temporal::FrameUBO_t) == 544 || sizeof(temporal::FrameUBO_t) == 560, "FrameUBO mirror");
}  // namespace compenv

namespace {
struct ImageDesc {  // mirrored by ctypes in oracle/pyoracle.py
    void* data;
    int32_t width, height, fmt, binding;
};
template <class F> void run_grid(int W, int H, int nThreads, F&& f) {
    if (nThreads < 1) nThreads = 1;
    std::vector<std::thread> th;
    for (int t = 0; t < nThreads; t++)
        th.emplace_back([&, t] {
            for (int y = t; y < H; y += nThreads)
                for (int x = 0; x < W; x++) {
                    compenv::gl_GlobalInvocationID = glm::uvec3((uint32_t)x, (uint32_t)y, 0u);
                    f();
                }
        });
    for (auto& t : th) t.join();
}
}  // namespace

extern "C" {
// Runs one compute pass over a W x H grid. shader: "temporal" | "variance" | "atrous" | "taa" | "tonemap" | "sharpen".
// ubo: the reference FrameUBO bytes (at least the shader's block size; ignored by tonemap / sharpen), pc: the
// shader's push-constant block (16 bytes; null where the shader has none). images: bound by the shader's binding numbers.
int ref_comp_run(const char* shader, const void* ubo, size_t uboBytes, const void* pc, const void* imagesV, int nImages, int W, int H,
                 int nThreads) {
    using namespace compenv;
    const ImageDesc* images = (const ImageDesc*)imagesV;
    const std::string s(shader);
#define BLOK_BIND(NS)                                                                                     \
    for (int i = 0; i < nImages; i++) {                                                                   \
        image2D im; im.data = images[i].data; im.width = images[i].width; im.height = images[i].height;   \
        im.fmt = images[i].fmt;                                                                           \
        if (!NS::bind_image(images[i].binding, im)) return -2;                                            \
    }
#define BLOK_UBO(NS, VAR)                                                                                 \
    {                                                                                                     \
        std::memset(&NS::VAR, 0, sizeof(NS::VAR));                                                        \
        std::memcpy(&NS::VAR, ubo, uboBytes < sizeof(NS::VAR) ? uboBytes : sizeof(NS::VAR));              \
    }
    if (s == "temporal") { BLOK_BIND(temporal) BLOK_UBO(temporal, frame) run_grid(W, H, nThreads, [] { temporal::shader_main(); }); }
    else if (s == "variance") { BLOK_BIND(variance) BLOK_UBO(variance, frame) run_grid(W, H, nThreads, [] { variance::shader_main(); }); }
    else if (s == "atrous") {
        BLOK_BIND(atrous) BLOK_UBO(atrous, frame) std::memcpy(&atrous::pc, pc, sizeof(atrous::pc));
        run_grid(W, H, nThreads, [] { atrous::shader_main(); });
    } else if (s == "taa") {
        BLOK_BIND(taa) BLOK_UBO(taa, ubo) std::memcpy(&taa::pc, pc, sizeof(taa::pc));
        run_grid(W, H, nThreads, [] { taa::shader_main(); });
    } else if (s == "tonemap") {
        BLOK_BIND(tonemap) std::memcpy(&tonemap::pc, pc, sizeof(tonemap::pc));
        run_grid(W, H, nThreads, [] { tonemap::shader_main(); });
    } else if (s == "sharpen") {
        BLOK_BIND(sharpen) std::memcpy(&sharpen::pc, pc, sizeof(sharpen::pc));
        run_grid(W, H, nThreads, [] { sharpen::shader_main(); });
    } else return -1;
    return 0;
}
// float <-> half as pinned above (used by the Python side to build rgba16f inputs the same way).
void ref_float_to_half(const float* in, uint16_t* out, size_t n) { for (size_t i = 0; i < n; i++) out[i] = compenv::float_to_half(in[i]); }
void ref_half_to_float(const uint16_t* in, float* out, size_t n) { for (size_t i = 0; i < n; i++) out[i] = compenv::half_to_float(in[i]); }
}
