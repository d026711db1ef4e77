// emul_shim.h -- TEST INFRASTRUCTURE ONLY.  Lets g++ compile the product's device headers
// (csrc/device_math.cuh, traverse.cuh, shade.cuh, path.cuh) for the HOST so the CPU-only test tier can check the
// traversal / shading LOGIC against the oracle without a GPU.  It is never linked into librt_b200.so and is not a
// fallback: the product has no CPU path.  Rounding-exact intrinsics map to plain IEEE ops (build with
// -ffp-contract=off); fast-math intrinsics map to their libm counterparts (so only X=true results are bit-exact).
#pragma once
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstring>

struct EmuDim { int x; };
static thread_local EmuDim emu_threadIdx{0}, emu_blockDim{1};
#define threadIdx emu_threadIdx
#define blockDim emu_blockDim

static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline float __frcp_rn(float a) { return 1.0f / a; }
static inline float __fdividef(float a, float b) { return a / b; }
#define __expf(a) expf(a)
static inline float rsqrtf(float a) { return 1.0f / sqrtf(a); }
static inline void sincospif(float a, float* s, float* c) { *s = sinf(3.14159265358979f * a); *c = cosf(3.14159265358979f * a); }
static inline uint32_t __umulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
template <class T> static inline T __ldg(const T* p) { return *p; }
template <class T> static inline T tex2D(cudaTextureObject_t, float, float) { return T{}; }  // RT_TEX_GPU needs the texture unit
