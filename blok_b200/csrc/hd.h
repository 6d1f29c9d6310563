// Qualifier macros so that traverse.cuh / shade.cuh also compile as plain host C++ — used ONLY by
// tests/host_emul (a CPU build of the very same device functions, to check the DDA's logic against the oracle
// in a container without a GPU). The product library never runs this path: libblokb200.so contains device code only.
#pragma once
#if defined(__CUDACC__)
#define BLOKB_ASSERT(c) ((void)0)
#define BLOKB_HD __device__ __forceinline__
#define BLOKB_D __device__ __forceinline__
#else
#include <cassert>
#include <cmath>
#define BLOKB_ASSERT(c) assert(c)
#include <cstdint>
#include <cstring>
#define BLOKB_HD inline
#define BLOKB_D inline
static inline int __float2int_rd(float v) { return (int)std::floor(v); }
static inline int __clz(uint32_t v) { return v ? __builtin_clz(v) : 32; }
static inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
static inline int __popc(uint32_t v) { return __builtin_popcount(v); }
static inline uint32_t __float_as_uint(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
static inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
#include <vector_types.h>
#ifndef __VECTOR_FUNCTIONS_H__
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { uint4 v; v.x = x; v.y = y; v.z = z; v.w = w; return v; }
#endif
template <class T> static inline T __ldg(const T* p) { return *p; }
#endif
