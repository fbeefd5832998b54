// Shared helpers for the ditsep_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/ditsep_b200.h"

namespace ditsep {

void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return 1;
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

#define DITSEP_REQUIRE(cond, ...)        \
  do {                                   \
    if (!(cond)) {                       \
      ::ditsep::set_error(__VA_ARGS__);  \
      return 2;                          \
    }                                    \
  } while (0)

#define DITSEP_CUDA(call)                                                        \
  do {                                                                           \
    cudaError_t e__ = (call);                                                    \
    if (e__ != cudaSuccess) {                                                    \
      ::ditsep::set_error("%s failed: %s", #call, cudaGetErrorString(e__));      \
      return 3;                                                                  \
    }                                                                            \
  } while (0)

constexpr int kNumSMs = 148;

typedef __nv_bfloat16 bf16;

// ---- typed scalar / vector access -------------------------------------------------
template <typename T>
__device__ __forceinline__ float to_f(T v);
template <>
__device__ __forceinline__ float to_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f<bf16>(bf16 v) { return __bfloat162float(v); }

template <typename T>
__device__ __forceinline__ T from_f(float v);
template <>
__device__ __forceinline__ float from_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// A "pack" is 8 consecutive channels: 32 B for f32 (two float4), 16 B for bf16 (one uint4).
template <typename T>
struct Pack8;
template <>
struct Pack8<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
    v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};
template <>
struct Pack8<bf16> {
  static __device__ __forceinline__ void load(const bf16* p, float (&v)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x;
      v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void store(bf16* p, const float (&v)[8]) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = u;
  }
};

// SiLU in exactly five instructions (FMUL, MUFU.EX2, FADD, MUFU.RCP, FMUL): explicit .ftz approximations,
// because `x / (1 + __expf(-x))` compiles to an IEEE division (GroupNorm-apply was issue-bound at 22
// instructions per element, r1e ncu) and even __fdividef / __expf carry denormal-range fix-ups without -ftz.
// ex2.approx and rcp.approx are good to 2 ulp; x -> -inf gives x * rcp(inf) = -0, x -> +inf gives x.
__device__ __forceinline__ float silu_f(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * -1.4426950408889634f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return x * r;
}

// Programmatic dependent launch, primary side: lets a dependent kernel launched with the programmatic-stream-
// serialization attribute (the conv tile kernel) be scheduled as this grid's CTAs retire.  A no-op otherwise.
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Secondary side: nothing below this call may run before the previous kernel in the stream has completed and its
// writes are visible.  A no-op for a kernel launched without the attribute.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Host side: launch with programmatic stream serialization (the kernel must call griddep_wait() before it touches
// global memory).  DITSEP_PDL=0 launches plainly.
// DITSEP_PDL: 0 = plain launches, 1 (default) = conv and GroupNorm kernels, 2 = FIR, gn_partial and attention too
// (measured on one box, batch 32: 10.91 / 10.85 / 10.88 ms per score step -- level 2 is no gain, so it stays opt-in).
inline int pdl_level() {
  static const int lvl = getenv("DITSEP_PDL") ? atoi(getenv("DITSEP_PDL")) : 1;
  return lvl;
}
inline bool pdl_enabled() { return pdl_level() >= 1; }
template <int kLevel = 1, typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_level() >= kLevel ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// SiLU with ONE MUFU: x * sigmoid(x) = h + h * tanh(h), h = x / 2 (FMUL, MUFU.TANH, FFMA).  tanh.approx.f32 is good to
// ~2^-11 relative, i.e. an absolute error of |h| * 5e-4 on the result: below bf16 rounding for positive outputs, up to a
// few bf16 ulps on the small negative tail.  Used only where the result is rounded to bf16 anyway AND the MUFU pipe is
// the bottleneck (the conv kernel's in-shared-memory GroupNorm transform; the bf16 FIR kernels, which activate every
// input element two (down) or three (up) times).
__device__ __forceinline__ float silu_tanh_f(float x) {
  const float h = 0.5f * x;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- order-independent fixed-point sums (GroupNorm statistics) --------------------------------------------------
// A sum is kept as two 64-bit integers {integer part, fraction in units of 2^-32}; a float v is split EXACTLY
// (floorf(v) and v - floorf(v) are exact in fp32; bits below 2^-32 of the fraction are dropped) and both parts are
// added with integer atomics.  Integer addition commutes, so the total does not depend on the order in which CTAs
// arrive: statistics are bit-reproducible and batch-invariant, with no per-CTA partial buffers and no finalize pass
// over hundreds of partials.  Range |sum| < 2^62, resolution 2^-32 absolute per addend.
__device__ __forceinline__ void fx_add(long long* dst, float v) {
  const float fl = floorf(v);
  const float fr = v - fl;  // in [0, 1), exact
  atomicAdd(reinterpret_cast<unsigned long long*>(dst), (unsigned long long)(long long)fl);
  atomicAdd(reinterpret_cast<unsigned long long*>(dst + 1), (unsigned long long)(fr * 4294967296.0f));
}
__device__ __forceinline__ double fx_value(const long long* p) {
  return (double)p[0] + (double)p[1] * (1.0 / 4294967296.0);
}

inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

}  // namespace ditsep
