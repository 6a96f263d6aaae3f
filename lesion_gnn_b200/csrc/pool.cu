// (3) global pooling = segment reduce over the sorted `batch` vector (global_mean_pool / global_max_pool /
// global_add_pool; reference models/gat.py:56, gin.py:33, pointnet.py:38).
//
// PyG does scatter_add_ / scatter_reduce_ with atomics on B*F addresses (n-bar rows fold into one).  Because
// collation keeps `batch` sorted, a graph is a contiguous [n_g, F] slab: one CTA streams the slab with 128-bit
// loads and reduces it in registers + shared memory -- no atomics, deterministic, sum and max (and the arg
// index the max backward needs) in the same pass.  Long graphs are split over gridDim.y CTAs with a small
// deterministic second stage.
#include "common.cuh"
#include <float.h>

namespace lgk {
namespace {

constexpr int kThreads = 256;
constexpr int kMaxSplits = 1024;

int choose_splits(int64_t N, int64_t B) {
  if (B <= 0 || N <= 0) return 1;
  int64_t want = ceil_div(int64_t(kNumSMs) * 4, B);  // enough CTAs to fill the machine
  int64_t cap = (N / B) / 64;                        // but at least 64 rows per CTA
  int64_t s = want < cap ? want : cap;
  if (s < 1) s = 1;
  if (s > kMaxSplits) s = kMaxSplits;
  return int(s);
}

struct PoolArgs {
  const int* ptr;
  int64_t n_graphs;
  int feat;
  int splits;
  int mean;
  // forward
  float* part_sum;  // [B, S, F] when splits > 1
  float* part_max;
  int* part_arg;
  int* arg;         // [B, F]
  int vec_ok;       // backward: F % 4 == 0 and every pointer 16-byte aligned
};

__device__ __forceinline__ void slice_of(const PoolArgs& a, int g, int sp, int& r0, int& r1, int& n) {
  const int start = a.ptr[g], end = a.ptr[g + 1];
  n = end - start;
  const int rps = (n + a.splits - 1) / a.splits;
  r0 = min(end, start + sp * rps);
  r1 = min(end, r0 + rps);
}

__device__ __forceinline__ void take_max(float& best, int& barg, float v, int r) {
  if (v > best || (v == best && r < barg)) { best = v; barg = r; }
}

// thread t owns feature(s) f = t, t + 256, ...  (scalar) -- or a float4 of features (vector path below)
template <typename T, bool DO_SUM, bool DO_MAX>
__global__ void __launch_bounds__(kThreads) pool_fwd_scalar_kernel(PoolArgs a, const T* __restrict__ x, T* __restrict__ out_sum,
                                                                   T* __restrict__ out_max) {
  const int g = blockIdx.x, sp = blockIdx.y, F = a.feat;
  int r0, r1, n;
  slice_of(a, g, sp, r0, r1, n);
  for (int f = threadIdx.x; f < F; f += kThreads) {
    float s = 0.f, best = -FLT_MAX;
    int barg = INT_MAX;
#pragma unroll 4
    for (int r = r0; r < r1; ++r) {
      const float v = Vec4<T>::load1(x + int64_t(r) * F, f);
      if (DO_SUM) s += v;
      if (DO_MAX) take_max(best, barg, v, r);
    }
    if (a.splits > 1) {
      const int64_t o = (int64_t(g) * a.splits + sp) * F + f;
      if (DO_SUM) a.part_sum[o] = s;
      if (DO_MAX) { a.part_max[o] = best; a.part_arg[o] = barg; }
    } else {
      const int64_t o = int64_t(g) * F + f;
      if (DO_SUM) Vec4<T>::store1(out_sum, o, a.mean ? s / float(max(n, 1)) : s);
      if (DO_MAX) {
        Vec4<T>::store1(out_max, o, n > 0 ? best : 0.f);
        if (a.arg) a.arg[o] = n > 0 ? barg : -1;
      }
    }
  }
}

// vector path: F % 4 == 0, F <= 512.  LPR lanes cover a row, 256/LPR rows are read per CTA step.
template <typename T, int LPR, int VPL, bool DO_SUM, bool DO_MAX>
__global__ void __launch_bounds__(kThreads) pool_fwd_vec_kernel(PoolArgs a, const T* __restrict__ x, T* __restrict__ out_sum,
                                                                T* __restrict__ out_max) {
  extern __shared__ __align__(16) float smem[];
  constexpr int NG = kThreads / LPR;
  const int g = blockIdx.x, sp = blockIdx.y, F = a.feat, F4 = F >> 2;
  const int gl = threadIdx.x & (LPR - 1), gi = threadIdx.x / LPR;
  int r0, r1, n;
  slice_of(a, g, sp, r0, r1, n);

  float4 s[VPL], best[VPL];
  int4 barg[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    s[v] = f4_zero();
    best[v] = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX);
    barg[v] = make_int4(INT_MAX, INT_MAX, INT_MAX, INT_MAX);
  }
#pragma unroll 4
  for (int r = r0 + gi; r < r1; r += NG) {
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int vec = gl + v * LPR;
      if (vec < F4) {
        const float4 val = Vec4<T>::stream(x + int64_t(r) * F, vec * 4);
        if (DO_SUM) { s[v].x += val.x; s[v].y += val.y; s[v].z += val.z; s[v].w += val.w; }
        if (DO_MAX) {
          // rows arrive in ascending order per group, so strict > keeps the first occurrence
          if (val.x > best[v].x) { best[v].x = val.x; barg[v].x = r; }
          if (val.y > best[v].y) { best[v].y = val.y; barg[v].y = r; }
          if (val.z > best[v].z) { best[v].z = val.z; barg[v].z = r; }
          if (val.w > best[v].w) { best[v].w = val.w; barg[v].w = r; }
        }
      }
    }
  }
  // cross-group reduction through shared memory: [NG][F] per quantity
  float* s_sum = smem;
  float* s_max = s_sum + (DO_SUM ? NG * F : 0);
  int* s_arg = reinterpret_cast<int*>(s_max + (DO_MAX ? NG * F : 0));
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    const int vec = gl + v * LPR;
    if (vec < F4) {
      if (DO_SUM) *reinterpret_cast<float4*>(s_sum + gi * F + vec * 4) = s[v];
      if (DO_MAX) {
        *reinterpret_cast<float4*>(s_max + gi * F + vec * 4) = best[v];
        *reinterpret_cast<int4*>(s_arg + gi * F + vec * 4) = barg[v];
      }
    }
  }
  __syncthreads();
  for (int f = threadIdx.x; f < F; f += kThreads) {
    float ts = 0.f, tb = -FLT_MAX;
    int ta = INT_MAX;
#pragma unroll 4
    for (int k = 0; k < NG; ++k) {
      if (DO_SUM) ts += s_sum[k * F + f];
      if (DO_MAX) take_max(tb, ta, s_max[k * F + f], s_arg[k * F + f]);
    }
    if (a.splits > 1) {
      const int64_t o = (int64_t(g) * a.splits + sp) * F + f;
      if (DO_SUM) a.part_sum[o] = ts;
      if (DO_MAX) { a.part_max[o] = tb; a.part_arg[o] = ta; }
    } else {
      const int64_t o = int64_t(g) * F + f;
      if (DO_SUM) Vec4<T>::store1(out_sum, o, a.mean ? ts / float(max(n, 1)) : ts);
      if (DO_MAX) {
        Vec4<T>::store1(out_max, o, n > 0 ? tb : 0.f);
        if (a.arg) a.arg[o] = n > 0 ? ta : -1;
      }
    }
  }
}

template <typename T, bool DO_SUM, bool DO_MAX>
__global__ void __launch_bounds__(kThreads) pool_combine_kernel(PoolArgs a, T* __restrict__ out_sum, T* __restrict__ out_max) {
  const int g = blockIdx.x, F = a.feat;
  const int n = a.ptr[g + 1] - a.ptr[g];
  for (int f = threadIdx.x; f < F; f += kThreads) {
    float ts = 0.f, tb = -FLT_MAX;
    int ta = INT_MAX;
    for (int sp = 0; sp < a.splits; ++sp) {
      const int64_t o = (int64_t(g) * a.splits + sp) * F + f;
      if (DO_SUM) ts += a.part_sum[o];
      if (DO_MAX) take_max(tb, ta, a.part_max[o], a.part_arg[o]);
    }
    const int64_t o = int64_t(g) * F + f;
    if (DO_SUM) Vec4<T>::store1(out_sum, o, a.mean ? ts / float(max(n, 1)) : ts);
    if (DO_MAX) {
      Vec4<T>::store1(out_max, o, n > 0 ? tb : 0.f);
      if (a.arg) a.arg[o] = n > 0 ? ta : -1;
    }
  }
}

// backward: dx[r,:] = dsum[g,:] * scale + (arg[g,f] == r) * dmax[g,f]
template <typename T>
__global__ void __launch_bounds__(kThreads) pool_bwd_kernel(PoolArgs a, const T* __restrict__ dsum, const T* __restrict__ dmax,
                                                            T* __restrict__ dx) {
  const int g = blockIdx.x, sp = blockIdx.y, F = a.feat;
  int r0, r1, n;
  slice_of(a, g, sp, r0, r1, n);
  const float scale = a.mean ? 1.f / float(max(n, 1)) : 1.f;
  if (a.vec_ok) {
    const int F4 = F >> 2;
    if (kThreads % F4 == 0) {
      // the usual widths (F4 divides the CTA): a thread keeps ONE column group for all its rows, so the graph's gradient
      // row, arg-max row and scale are loaded once and the row loop is compare / add / store (the general loop below divides
      // a 64-bit counter per element: 6.8 us for the 11 k x 64 headline batch, where 3 MB move)
      const int vec = threadIdx.x % F4, rstep = kThreads / F4;
      float4 base = f4_zero(), dm = f4_zero();
      int4 ar = make_int4(-1, -1, -1, -1);
      if (dsum) {
        base = Vec4<T>::gather(dsum + int64_t(g) * F, vec * 4);
        base.x *= scale; base.y *= scale; base.z *= scale; base.w *= scale;
      }
      if (dmax) {
        ar = __ldg(reinterpret_cast<const int4*>(a.arg + int64_t(g) * F) + vec);
        dm = Vec4<T>::gather(dmax + int64_t(g) * F, vec * 4);
      }
      for (int r = r0 + threadIdx.x / F4; r < r1; r += rstep) {
        float4 o = base;
        if (ar.x == r) o.x += dm.x;
        if (ar.y == r) o.y += dm.y;
        if (ar.z == r) o.z += dm.z;
        if (ar.w == r) o.w += dm.w;
        Vec4<T>::store(dx + int64_t(r) * F, vec * 4, o);
      }
      return;
    }
    // thread owns one float4 column group and walks down the rows; consecutive threads -> consecutive vectors
    const int64_t total = int64_t(r1 - r0) * F4;
    for (int64_t t = threadIdx.x; t < total; t += kThreads) {
      const int r = r0 + int(t / F4), vec = int(t % F4);
      float4 o = f4_zero();
      if (dsum) {
        o = Vec4<T>::gather(dsum + int64_t(g) * F, vec * 4);
        o.x *= scale; o.y *= scale; o.z *= scale; o.w *= scale;
      }
      if (dmax) {
        const int4 ar = __ldg(reinterpret_cast<const int4*>(a.arg + int64_t(g) * F) + vec);
        const float4 dm = Vec4<T>::gather(dmax + int64_t(g) * F, vec * 4);
        if (ar.x == r) o.x += dm.x;
        if (ar.y == r) o.y += dm.y;
        if (ar.z == r) o.z += dm.z;
        if (ar.w == r) o.w += dm.w;
      }
      Vec4<T>::store(dx + int64_t(r) * F, vec * 4, o);
    }
  } else {
    const int64_t total = int64_t(r1 - r0) * F;
    for (int64_t t = threadIdx.x; t < total; t += kThreads) {
      const int r = r0 + int(t / F), f = int(t % F);
      float o = 0.f;
      if (dsum) o = Vec4<T>::load1(dsum + int64_t(g) * F, f) * scale;
      if (dmax && __ldg(a.arg + int64_t(g) * F + f) == r) o += Vec4<T>::load1(dmax + int64_t(g) * F, f);
      Vec4<T>::store1(dx + int64_t(r) * F, f, o);
    }
  }
}

template <typename T, bool DO_SUM, bool DO_MAX>
int launch_fwd(const PoolArgs& a, const void* x_, void* out_sum_, void* out_max_, cudaStream_t stream) {
  const T* x = static_cast<const T*>(x_);
  T* out_sum = static_cast<T*>(out_sum_);
  T* out_max = static_cast<T*>(out_max_);
  const int F = a.feat;
  dim3 grid(unsigned(a.n_graphs), unsigned(a.splits));
  const bool aligned = (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  if (F % 4 == 0 && F <= 512 && aligned) {
    const int F4 = F / 4;
    const int quantities = (DO_SUM ? 1 : 0) + (DO_MAX ? 2 : 0);
#define LGK_POOL_CASE(LPR, VPL)                                                                   \
  {                                                                                               \
    size_t smem = size_t(kThreads / (LPR)) * F * 4 * quantities;                                  \
    pool_fwd_vec_kernel<T, LPR, VPL, DO_SUM, DO_MAX><<<grid, kThreads, smem, stream>>>(a, x, out_sum, out_max); \
  }
    if (F4 <= 1) LGK_POOL_CASE(1, 1)
    else if (F4 <= 2) LGK_POOL_CASE(2, 1)
    else if (F4 <= 4) LGK_POOL_CASE(4, 1)
    else if (F4 <= 8) LGK_POOL_CASE(8, 1)
    else if (F4 <= 16) LGK_POOL_CASE(16, 1)
    else if (F4 <= 32) LGK_POOL_CASE(32, 1)
    else if (F4 <= 64) LGK_POOL_CASE(32, 2)
    else LGK_POOL_CASE(32, 4)
#undef LGK_POOL_CASE
  } else {
    pool_fwd_scalar_kernel<T, DO_SUM, DO_MAX><<<grid, kThreads, 0, stream>>>(a, x, out_sum, out_max);
  }
  LGK_CHECK_LAUNCH("pool_fwd_kernel");
  if (a.splits > 1) {
    pool_combine_kernel<T, DO_SUM, DO_MAX><<<unsigned(a.n_graphs), kThreads, 0, stream>>>(a, out_sum, out_max);
    LGK_CHECK_LAUNCH("pool_combine_kernel");
  }
  return LGK_OK;
}

template <typename T>
int dispatch_fwd(const PoolArgs& a, const void* x, void* out_sum, void* out_max, cudaStream_t stream) {
  if (out_sum && out_max) return launch_fwd<T, true, true>(a, x, out_sum, out_max, stream);
  if (out_sum) return launch_fwd<T, true, false>(a, x, out_sum, out_max, stream);
  return launch_fwd<T, false, true>(a, x, out_sum, out_max, stream);
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int64_t lgk_pool_workspace_bytes(int64_t N, int64_t B, int64_t F) {
  if (N <= 0 || B <= 0 || F <= 0) return 0;
  int s = choose_splits(N, B);
  if (s <= 1) return 0;
  return align_up(B * s * F * 4, 256) * 3;
}

extern "C" int lgk_pool_fwd(const void* x, const int32_t* ptr, void* out_sum, int mean, void* out_max, int32_t* arg,
                            int64_t N, int64_t B, int64_t F, void* workspace, int64_t workspace_bytes, int dtype,
                            void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(N >= 0 && B >= 0 && F >= 0, LGK_ERR_BAD_ARG, "lgk_pool_fwd: negative size");
  if (B == 0 || F == 0) return LGK_OK;
  LGK_REQUIRE(ptr && (out_sum || out_max) && (N == 0 || x), LGK_ERR_BAD_ARG, "lgk_pool_fwd: null pointer");
  LGK_REQUIRE(B < (1 << 30) && F < (1 << 24), LGK_ERR_UNSUPPORTED, "lgk_pool_fwd: size out of range");
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_BAD_ARG, "lgk_pool_fwd: bad dtype");
  PoolArgs a{};
  a.ptr = ptr; a.n_graphs = B; a.feat = int(F); a.mean = mean; a.arg = arg;
  a.splits = choose_splits(N, B);
  if (a.splits > 1) {
    const int64_t need = lgk_pool_workspace_bytes(N, B, F);
    LGK_REQUIRE(workspace && workspace_bytes >= need, LGK_ERR_WORKSPACE, "lgk_pool_fwd: workspace %lld < %lld bytes",
                (long long)workspace_bytes, (long long)need);
    const int64_t chunk = align_up(B * a.splits * F * 4, 256);
    char* w = static_cast<char*>(workspace);
    a.part_sum = reinterpret_cast<float*>(w);
    a.part_max = reinterpret_cast<float*>(w + chunk);
    a.part_arg = reinterpret_cast<int*>(w + 2 * chunk);
  }
  return dtype == LGK_F32 ? dispatch_fwd<float>(a, x, out_sum, out_max, stream)
                          : dispatch_fwd<__nv_bfloat16>(a, x, out_sum, out_max, stream);
}

extern "C" int lgk_pool_bwd(const void* dsum, int mean, const void* dmax, const int32_t* arg, const int32_t* ptr, void* dx,
                            int64_t N, int64_t B, int64_t F, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(N >= 0 && B >= 0 && F >= 0, LGK_ERR_BAD_ARG, "lgk_pool_bwd: negative size");
  if (N == 0 || B == 0 || F == 0) return LGK_OK;
  LGK_REQUIRE(ptr && dx && (dsum || dmax) && (!dmax || arg), LGK_ERR_BAD_ARG, "lgk_pool_bwd: null pointer");
  PoolArgs a{};
  a.ptr = ptr; a.n_graphs = B; a.feat = int(F); a.mean = mean; a.arg = const_cast<int*>(arg);
  a.splits = choose_splits(N, B);
  a.vec_ok = (F % 4 == 0) && ((reinterpret_cast<uintptr_t>(dsum) | reinterpret_cast<uintptr_t>(dmax) |
                               reinterpret_cast<uintptr_t>(arg) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0;
  dim3 grid(unsigned(B), unsigned(a.splits));
  if (dtype == LGK_F32)
    pool_bwd_kernel<float><<<grid, kThreads, 0, stream>>>(a, static_cast<const float*>(dsum), static_cast<const float*>(dmax),
                                                          static_cast<float*>(dx));
  else if (dtype == LGK_BF16)
    pool_bwd_kernel<__nv_bfloat16><<<grid, kThreads, 0, stream>>>(a, static_cast<const __nv_bfloat16*>(dsum),
                                                                  static_cast<const __nv_bfloat16*>(dmax),
                                                                  static_cast<__nv_bfloat16*>(dx));
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_pool_bwd: bad dtype");
  LGK_CHECK_LAUNCH("pool_bwd_kernel");
  return LGK_OK;
}
