// Shared device/host helpers for liblgk (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/lgk.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "liblgk is written for sm_100a (B200) only"
#endif

namespace lgk {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// ---- host-side error plumbing -------------------------------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define LGK_REQUIRE(cond, code, ...)      \
  do {                                    \
    if (!(cond)) {                        \
      ::lgk::set_error(__VA_ARGS__);      \
      return (code);                      \
    }                                     \
  } while (0)

// checked after every launch; never synchronises
#define LGK_CHECK_LAUNCH(name)                                                       \
  do {                                                                               \
    cudaError_t e__ = cudaGetLastError();                                            \
    if (e__ != cudaSuccess) {                                                        \
      ::lgk::set_error("%s: launch failed: %s", (name), cudaGetErrorString(e__));    \
      return LGK_ERR_LAUNCH;                                                         \
    }                                                                                \
    ::lgk::count_launch();                                                           \
  } while (0)

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t align_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

// ---- device helpers -----------------------------------------------------------------------------
#ifdef __CUDACC__

// streaming (read-once) loads: do not allocate in L1 so the gathered feature rows keep it
__device__ __forceinline__ int ld_stream_i32(const int* p) {
  int v;
  asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_stream_f32(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
// gathered rows: read-only path, L1-allocating (neighbouring rows of a CTA share sources)
__device__ __forceinline__ float4 ld_gather_f4(const float4* p) { return __ldg(p); }
__device__ __forceinline__ uint2 ld_gather_u2(const uint2* p) { return __ldg(p); }

// predicated gathers: the load is issued only when `on` (no L1 wavefront otherwise) and the result is 0 when off.
// Written as volatile asm so that ptxas keeps a run of them back to back (independent loads in flight together)
// instead of sinking each one next to its consumer.
__device__ __forceinline__ float4 ld_gather_f4_pred(const float4* p, bool on) {
  float4 v;
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
      "mov.f32 %0, 0f00000000;\n\tmov.f32 %1, 0f00000000;\n\tmov.f32 %2, 0f00000000;\n\tmov.f32 %3, 0f00000000;\n\t"
      "@p ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];\n\t}"
      : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
      : "l"(p), "r"(int(on)));
  return v;
}
__device__ __forceinline__ uint2 ld_gather_u2_pred(const uint2* p, bool on) {
  uint2 v;
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\tmov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\t"
      "@p ld.global.nc.v2.u32 {%0,%1}, [%2];\n\t}"
      : "=r"(v.x), "=r"(v.y)
      : "l"(p), "r"(int(on)));
  return v;
}

// streaming stores (written once, consumed by a later kernel): evict-first
__device__ __forceinline__ void st_stream_f4(float4* p, float4 v) { __stcs(p, v); }
__device__ __forceinline__ void st_stream_u2(uint2* p, uint2 v) { __stcs(p, v); }
__device__ __forceinline__ void st_stream_i4(int4* p, int4 v) { __stcs(p, v); }

__device__ __forceinline__ float4 bf16x4_to_f4(uint2 u) {
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&u.x);
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&u.y);
  float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
  return make_float4(fa.x, fa.y, fb.x, fb.y);
}
__device__ __forceinline__ uint2 f4_to_bf16x4(float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
  __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  return u;
}

// A "row" of features is handled 4 elements per lane.  T = float (16 B per lane) or bf16 (8 B per lane).
template <typename T>
struct Vec4;
template <>
struct Vec4<float> {
  static __device__ __forceinline__ float4 gather(const float* base, int64_t elem) {
    return ld_gather_f4(reinterpret_cast<const float4*>(base + elem));
  }
  static __device__ __forceinline__ float4 gather_if(const float* base, int64_t elem, bool on) {
    return ld_gather_f4_pred(reinterpret_cast<const float4*>(base + elem), on);
  }
  static __device__ __forceinline__ float4 stream(const float* base, int64_t elem) {
    return ld_stream_f4(reinterpret_cast<const float4*>(base + elem));
  }
  static __device__ __forceinline__ void store(float* base, int64_t elem, float4 v) {
    st_stream_f4(reinterpret_cast<float4*>(base + elem), v);
  }
  static __device__ __forceinline__ float load1(const float* base, int64_t elem) { return __ldg(base + elem); }
  static __device__ __forceinline__ void store1(float* base, int64_t elem, float v) { base[elem] = v; }
};
template <>
struct Vec4<__nv_bfloat16> {
  static __device__ __forceinline__ float4 gather(const __nv_bfloat16* base, int64_t elem) {
    return bf16x4_to_f4(ld_gather_u2(reinterpret_cast<const uint2*>(base + elem)));
  }
  static __device__ __forceinline__ float4 gather_if(const __nv_bfloat16* base, int64_t elem, bool on) {
    return bf16x4_to_f4(ld_gather_u2_pred(reinterpret_cast<const uint2*>(base + elem), on));
  }
  static __device__ __forceinline__ float4 stream(const __nv_bfloat16* base, int64_t elem) {
    return bf16x4_to_f4(ld_gather_u2(reinterpret_cast<const uint2*>(base + elem)));
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* base, int64_t elem, float4 v) {
    st_stream_u2(reinterpret_cast<uint2*>(base + elem), f4_to_bf16x4(v));
  }
  static __device__ __forceinline__ float load1(const __nv_bfloat16* base, int64_t elem) {
    return __bfloat162float(base[elem]);
  }
  static __device__ __forceinline__ void store1(__nv_bfloat16* base, int64_t elem, float v) {
    base[elem] = __float2bfloat16_rn(v);
  }
};

__device__ __forceinline__ float4 f4_zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ void f4_fma(float4& acc, float s, const float4& v) {
  acc.x = fmaf(s, v.x, acc.x);
  acc.y = fmaf(s, v.y, acc.y);
  acc.z = fmaf(s, v.z, acc.z);
  acc.w = fmaf(s, v.w, acc.w);
}
__device__ __forceinline__ float f4_dot(const float4& a, const float4& b) {
  return fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, a.w * b.w)));
}
__device__ __forceinline__ float elu1(float v) { return v > 0.f ? v : expm1f(v); }

// Philox-free counter RNG for attention dropout: 2 rounds of a 64-bit mix (splitmix64 finaliser) on
// (seed, slot, head) -> uniform in [0,1).  Forward and both backward passes evaluate the same function,
// so the mask never has to be stored.
__device__ __forceinline__ float uniform_hash(uint64_t seed, uint32_t slot, uint32_t head) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * ((uint64_t(slot) << 8) ^ uint64_t(head + 1));
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return float(uint32_t(z >> 40)) * (1.0f / 16777216.0f);
}

// attention-logit halves computed in the epilogue of GATConv's transform (lgk_linear_gat_fwd)
struct GatAlpha {
  const float* att_src;  // [heads * chan]
  const float* att_dst;
  float* a_src;          // [n, heads]
  float* a_dst;
  int heads, chan;
};

// Second stage of every slab reduction in the library (weight / bias / attention-vector gradients): out (+)= sum over
// slabs of partial[slab][t], t < count_a goes to out_a, the rest to out_b.  32 outputs x 8 slab lanes per CTA: a
// lane adds its slabs in ascending order, the 8 lane sums are combined in a fixed order -- deterministic, and the
// ~150 partials per output are read with 8 x 4 loads in flight instead of one dependent chain.
static __global__ void __launch_bounds__(256) slab_reduce_kernel(const float* __restrict__ partial, float* __restrict__ out_a,
                                                                 float* __restrict__ out_b, int64_t count_a, int64_t count_b,
                                                                 int slabs, int accumulate) {
  __shared__ float red[8][33];
  const int o = threadIdx.x & 31, sl = threadIdx.x >> 5;
  const int64_t t = int64_t(blockIdx.x) * 32 + o;
  const int64_t stride = count_a + count_b;
  float s0 = 0.f, s1 = 0.f;
  if (t < stride) {
    int k = sl;
    for (; k + 8 < slabs; k += 16) {
      s0 += __ldg(partial + int64_t(k) * stride + t);
      s1 += __ldg(partial + int64_t(k + 8) * stride + t);
    }
    if (k < slabs) s0 += __ldg(partial + int64_t(k) * stride + t);
  }
  red[sl][o] = s0 + s1;
  __syncthreads();
  if (sl == 0 && t < stride) {
    float s = red[0][o];
#pragma unroll
    for (int l = 1; l < 8; ++l) s += red[l][o];
    float* dst = t < count_a ? (out_a ? out_a + t : nullptr) : (out_b ? out_b + (t - count_a) : nullptr);
    if (dst) {
      if (accumulate) s += *dst;
      *dst = s;
    }
  }
}
static inline unsigned slab_reduce_grid(int64_t count) { return unsigned((count + 31) / 32); }

#endif  // __CUDACC__
}  // namespace lgk
