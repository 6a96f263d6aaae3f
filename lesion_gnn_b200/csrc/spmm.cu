// (2)+(4) fused gather + segment-reduce aggregation (SpMM) over a CSR (forward) or CSC (backward).
//
// Replaces index_select + mul + scatter(sum|mean|max) of MessagePassing.propagate (reference gin.py:31,
// drgnet.py:55) and torch_sparse.matmul on the adj_t path.  HBM/L2-bound: no tensor cores.
//
// Mapping: a feature row is split into 128-bit pieces, one piece per lane; a row needs LPR lanes
// (F/4 rounded up to a power of two, at most 32) so a warp reduces 32/LPR rows at once ("warp per row" for
// F >= 128, sub-warp groups below).  Each group loads LPR neighbour indices (+ weights) of its row with ONE
// coalesced streaming load, then broadcasts them with warp shuffles while every lane gathers its 16-byte
// piece of the neighbour's row through the read-only, L1-allocating path.  Branch-free inner loop (skipped
// entries become weight-0 reads of the own row) so that the unrolled gathers are all in flight together.
#include "common.cuh"
#include <float.h>

namespace lgk {
namespace {

constexpr int kThreads = 256;

struct SpmmArgs {
  const int* ptr;
  const int* idx;
  const int* emap;
  const float* ew;
  const float* pre;
  const float* post;
  const float* selfw;
  const float* bias;
  int* arg_out;
  float self_fill;
  int self_mode;
  int64_t n_rows;
  int feat;
};

template <int LPR>
__device__ __forceinline__ unsigned group_mask() {
  if constexpr (LPR == 32) {
    return 0xffffffffu;
  } else {
    unsigned lane = threadIdx.x & 31;
    return ((1u << LPR) - 1u) << (lane & ~(LPR - 1));
  }
}

// ---- vectorised path: feat % 4 == 0 ------------------------------------------------------------------
template <typename T, int LPR, int VPL, int REDUCE>
__global__ void __launch_bounds__(kThreads) spmm_vec_kernel(SpmmArgs a, const T* __restrict__ x, T* __restrict__ out) {
  constexpr int RPW = 32 / LPR;
  const int gl = threadIdx.x & (LPR - 1);  // lane inside the group
  const int64_t row = (int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5)) * RPW + ((threadIdx.x & 31) / LPR);
  if (row >= a.n_rows) return;
  const unsigned mask = group_mask<LPR>();
  const int F = a.feat, F4 = F >> 2;
  const int start = a.ptr[row], end = a.ptr[row + 1];

  float4 acc[VPL];
  int4 arg[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    acc[v] = (REDUCE == LGK_MAX) ? make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX) : f4_zero();
    arg[v] = make_int4(-1, -1, -1, -1);
  }

  for (int base = start; base < end; base += LPR) {
    const int p = base + gl;
    int my_j = int(row);
    float my_w = 0.f;
    if (p < end) {
      int j = ld_stream_i32(a.idx + p);
      float w = 1.f;
      if (a.ew) w = __ldg(a.ew + (a.emap ? ld_stream_i32(a.emap + p) : p));
      if (a.pre) w *= __ldg(a.pre + j);
      if (a.self_mode == LGK_SELF_REPLACE && j == int(row)) w = 0.f;  // stored loop: dropped
      my_j = j;
      my_w = w;
    }
    const int cnt = min(LPR, end - base);
#pragma unroll 4
    for (int u = 0; u < cnt; ++u) {
      const int j = __shfl_sync(mask, my_j, u, LPR);
      const float w = __shfl_sync(mask, my_w, u, LPR);
      const T* xr = x + int64_t(j) * F;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const int vec = gl + v * LPR;
        if (vec < F4) {
          float4 val = Vec4<T>::gather(xr, vec * 4);
          if (REDUCE == LGK_MAX) {
            const int slot = base + u;
            if (val.x > acc[v].x) { acc[v].x = val.x; arg[v].x = slot; }
            if (val.y > acc[v].y) { acc[v].y = val.y; arg[v].y = slot; }
            if (val.z > acc[v].z) { acc[v].z = val.z; arg[v].z = slot; }
            if (val.w > acc[v].w) { acc[v].w = val.w; arg[v].w = slot; }
          } else {
            f4_fma(acc[v], w, val);
          }
        }
      }
    }
  }

  float scale = 1.f;
  if (REDUCE == LGK_MEAN) scale = 1.f / float(max(end - start, 1));
  if (a.post) scale *= __ldg(a.post + row);
  float wself = 0.f;
  if (REDUCE != LGK_MAX && a.self_mode != LGK_SELF_NONE) {
    wself = a.selfw ? __ldg(a.selfw + row) : a.self_fill;
    if (a.pre) wself *= __ldg(a.pre + row);
  }
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    const int vec = gl + v * LPR;
    if (vec < F4) {
      float4 r = acc[v];
      if (REDUCE == LGK_MAX) {
        if (arg[v].x < 0) r.x = 0.f;
        if (arg[v].y < 0) r.y = 0.f;
        if (arg[v].z < 0) r.z = 0.f;
        if (arg[v].w < 0) r.w = 0.f;
        if (a.arg_out) st_stream_i4(reinterpret_cast<int4*>(a.arg_out + row * F + vec * 4), arg[v]);
      } else {
        if (a.self_mode != LGK_SELF_NONE) f4_fma(r, wself, Vec4<T>::gather(x + row * F, vec * 4));
        r.x *= scale; r.y *= scale; r.z *= scale; r.w *= scale;
      }
      if (a.bias) {
        float4 b = __ldg(reinterpret_cast<const float4*>(a.bias) + vec);
        r.x += b.x; r.y += b.y; r.z += b.z; r.w += b.w;
      }
      Vec4<T>::store(out + row * F, vec * 4, r);
    }
  }
}

// ---- scalar path: any feat (1025-wide raw lesion features, DRGNet's 1-wide last layer, ...) ----------
// warp per row, lanes stride over the feature dimension in chunks of 32 * UNR
template <typename T, int REDUCE>
__global__ void __launch_bounds__(kThreads) spmm_scalar_kernel(SpmmArgs a, const T* __restrict__ x, T* __restrict__ out) {
  constexpr int UNR = 4;
  const int lane = threadIdx.x & 31;
  const int64_t row = int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5);
  if (row >= a.n_rows) return;
  const int F = a.feat;
  const int start = a.ptr[row], end = a.ptr[row + 1];
  float scale = 1.f;
  if (REDUCE == LGK_MEAN) scale = 1.f / float(max(end - start, 1));
  if (a.post) scale *= __ldg(a.post + row);
  float wself = 0.f;
  if (REDUCE != LGK_MAX && a.self_mode != LGK_SELF_NONE) {
    wself = a.selfw ? __ldg(a.selfw + row) : a.self_fill;
    if (a.pre) wself *= __ldg(a.pre + row);
  }
  for (int f0 = 0; f0 < F; f0 += 32 * UNR) {
    float acc[UNR];
    int arg[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) { acc[u] = (REDUCE == LGK_MAX) ? -FLT_MAX : 0.f; arg[u] = -1; }
    for (int base = start; base < end; base += 32) {
      const int p = base + lane;
      int my_j = int(row);
      float my_w = 0.f;
      if (p < end) {
        int j = ld_stream_i32(a.idx + p);
        float w = 1.f;
        if (a.ew) w = __ldg(a.ew + (a.emap ? ld_stream_i32(a.emap + p) : p));
        if (a.pre) w *= __ldg(a.pre + j);
        if (a.self_mode == LGK_SELF_REPLACE && j == int(row)) w = 0.f;
        my_j = j;
        my_w = w;
      }
      const int cnt = min(32, end - base);
      for (int t = 0; t < cnt; ++t) {
        const int j = __shfl_sync(0xffffffffu, my_j, t);
        const float w = __shfl_sync(0xffffffffu, my_w, t);
        const T* xr = x + int64_t(j) * F;
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          const int f = f0 + u * 32 + lane;
          if (f < F) {
            float val = Vec4<T>::load1(xr, f);
            if (REDUCE == LGK_MAX) {
              if (val > acc[u]) { acc[u] = val; arg[u] = base + t; }
            } else {
              acc[u] = fmaf(w, val, acc[u]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int f = f0 + u * 32 + lane;
      if (f < F) {
        float r = acc[u];
        if (REDUCE == LGK_MAX) {
          if (arg[u] < 0) r = 0.f;
          if (a.arg_out) a.arg_out[row * F + f] = arg[u];
        } else {
          if (a.self_mode != LGK_SELF_NONE) r = fmaf(wself, Vec4<T>::load1(x + row * F, f), r);
          r *= scale;
        }
        if (a.bias) r += __ldg(a.bias + f);
        Vec4<T>::store1(out + row * F, f, r);
      }
    }
  }
}

// ---- backward of max: CSC gather of dy where the forward's winning slot matches --------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) spmm_max_bwd_kernel(const int* __restrict__ colptr, const int* __restrict__ rowidx,
                                                                const int* __restrict__ csr2csc, const int* __restrict__ arg,
                                                                const T* __restrict__ dy, T* __restrict__ dx, int64_t n_rows, int F) {
  const int lane = threadIdx.x & 31;
  const int64_t j = int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5);
  if (j >= n_rows) return;
  const int start = colptr[j], end = colptr[j + 1];
  for (int f = lane; f < F; f += 32) {
    float acc = 0.f;
    for (int q = start; q < end; ++q) {
      const int i = __ldg(rowidx + q), slot = __ldg(csr2csc + q);
      if (__ldg(arg + int64_t(i) * F + f) == slot) acc += Vec4<T>::load1(dy + int64_t(i) * F, f);
    }
    Vec4<T>::store1(dx + j * F, f, acc);
  }
}

__global__ void __launch_bounds__(256) gcn_norm_kernel(const int* __restrict__ rowptr, const int* __restrict__ col,
                                                       const float* __restrict__ ew, float fill, int add_self_loops,
                                                       float* __restrict__ dinv, float* __restrict__ selfw_out, int64_t N) {
  int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int start = rowptr[i], end = rowptr[i + 1];
  float deg = 0.f, sw = fill;
  for (int p = start; p < end; ++p) {
    const int j = __ldg(col + p);
    const float w = ew ? __ldg(ew + p) : 1.f;
    if (add_self_loops && j == int(i)) {
      // stored loops sit next to each other in ascending original edge id: the last one wins
      if (ew) sw = w;
    } else {
      deg += w;
    }
  }
  if (add_self_loops) deg += sw;
  dinv[i] = deg == 0.f ? 0.f : 1.0f / sqrtf(deg);
  if (selfw_out) selfw_out[i] = sw;
}

template <typename T, int REDUCE>
int launch_spmm(const SpmmArgs& a, const void* x_, void* out_, cudaStream_t stream) {
  const T* x = static_cast<const T*>(x_);
  T* out = static_cast<T*>(out_);
  const int F = a.feat;
  const int warps = kThreads / 32;
  const bool aligned = ((reinterpret_cast<uintptr_t>(x_) | reinterpret_cast<uintptr_t>(out_) |
                         reinterpret_cast<uintptr_t>(a.bias) | reinterpret_cast<uintptr_t>(a.arg_out)) & 15) == 0;
  if (F % 4 == 0 && F <= 1024 && aligned) {
    const int F4 = F / 4;
#define LGK_SPMM_CASE(LPR, VPL)                                                                     \
  {                                                                                                 \
    const int64_t rows_per_cta = int64_t(warps) * (32 / (LPR));                                     \
    spmm_vec_kernel<T, LPR, VPL, REDUCE><<<unsigned(ceil_div(a.n_rows, rows_per_cta)), kThreads, 0, stream>>>(a, x, out); \
  }
    if (F4 <= 1) LGK_SPMM_CASE(1, 1)
    else if (F4 <= 2) LGK_SPMM_CASE(2, 1)
    else if (F4 <= 4) LGK_SPMM_CASE(4, 1)
    else if (F4 <= 8) LGK_SPMM_CASE(8, 1)
    else if (F4 <= 16) LGK_SPMM_CASE(16, 1)
    else if (F4 <= 32) LGK_SPMM_CASE(32, 1)
    else if (F4 <= 64) LGK_SPMM_CASE(32, 2)
    else if (F4 <= 128) LGK_SPMM_CASE(32, 4)
    else LGK_SPMM_CASE(32, 8)
#undef LGK_SPMM_CASE
  } else {
    spmm_scalar_kernel<T, REDUCE><<<unsigned(ceil_div(a.n_rows, warps)), kThreads, 0, stream>>>(a, x, out);
  }
  LGK_CHECK_LAUNCH("spmm_kernel");
  return LGK_OK;
}

template <typename T>
int dispatch_reduce(const SpmmArgs& a, int reduce, const void* x, void* out, cudaStream_t stream) {
  switch (reduce) {
    case LGK_SUM: return launch_spmm<T, LGK_SUM>(a, x, out, stream);
    case LGK_MEAN: return launch_spmm<T, LGK_MEAN>(a, x, out, stream);
    case LGK_MAX: return launch_spmm<T, LGK_MAX>(a, x, out, stream);
  }
  set_error("lgk_spmm: unknown reduce %d", reduce);
  return LGK_ERR_BAD_ARG;
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int lgk_spmm(const int32_t* ptr, const int32_t* idx, const int32_t* emap, const float* ew, const float* pre,
                        const float* post, const float* selfw, float self_fill, int self_mode, const void* x,
                        const float* bias, void* out, int32_t* arg_out, int64_t n_rows, int64_t feat, int reduce,
                        int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n_rows >= 0 && feat >= 0, LGK_ERR_BAD_ARG, "lgk_spmm: negative size");
  if (n_rows == 0 || feat == 0) return LGK_OK;
  LGK_REQUIRE(ptr && x && out, LGK_ERR_BAD_ARG, "lgk_spmm: null pointer");
  LGK_REQUIRE(feat < (1 << 30) && n_rows < (int64_t(1) << 31), LGK_ERR_UNSUPPORTED, "lgk_spmm: size out of range");
  LGK_REQUIRE(self_mode >= LGK_SELF_NONE && self_mode <= LGK_SELF_ADD, LGK_ERR_BAD_ARG, "lgk_spmm: bad self_mode");
  LGK_REQUIRE(reduce == LGK_SUM || self_mode == LGK_SELF_NONE, LGK_ERR_BAD_ARG,
              "lgk_spmm: self-loop handling is defined for the sum reduction only");
  LGK_REQUIRE(reduce != LGK_MAX || (!ew && !pre && !post), LGK_ERR_BAD_ARG, "lgk_spmm: max takes no weights");
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_BAD_ARG, "lgk_spmm: bad dtype %d", dtype);
  LGK_REQUIRE(x != out, LGK_ERR_BAD_ARG, "lgk_spmm: in-place aggregation is not possible");
  SpmmArgs a{ptr, idx, emap, ew, pre, post, selfw, bias, arg_out, self_fill, self_mode, n_rows, int(feat)};
  return dtype == LGK_F32 ? dispatch_reduce<float>(a, reduce, x, out, stream)
                          : dispatch_reduce<__nv_bfloat16>(a, reduce, x, out, stream);
}

extern "C" int lgk_spmm_max_bwd(const int32_t* colptr, const int32_t* row, const int32_t* csr2csc, const int32_t* arg,
                                const void* dy, void* dx, int64_t n_rows, int64_t feat, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n_rows >= 0 && feat >= 0, LGK_ERR_BAD_ARG, "lgk_spmm_max_bwd: negative size");
  if (n_rows == 0 || feat == 0) return LGK_OK;
  LGK_REQUIRE(colptr && arg && dy && dx, LGK_ERR_BAD_ARG, "lgk_spmm_max_bwd: null pointer");
  unsigned grid = unsigned(ceil_div(n_rows, kThreads / 32));
  if (dtype == LGK_F32)
    spmm_max_bwd_kernel<float><<<grid, kThreads, 0, stream>>>(colptr, row, csr2csc, arg, static_cast<const float*>(dy),
                                                              static_cast<float*>(dx), n_rows, int(feat));
  else if (dtype == LGK_BF16)
    spmm_max_bwd_kernel<__nv_bfloat16><<<grid, kThreads, 0, stream>>>(colptr, row, csr2csc, arg,
                                                                      static_cast<const __nv_bfloat16*>(dy),
                                                                      static_cast<__nv_bfloat16*>(dx), n_rows, int(feat));
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_spmm_max_bwd: bad dtype %d", dtype);
  LGK_CHECK_LAUNCH("spmm_max_bwd_kernel");
  return LGK_OK;
}

extern "C" int lgk_gcn_norm(const int32_t* rowptr, const int32_t* col, const float* ew, float fill, int add_self_loops,
                            float* dinv, float* selfw_out, int64_t num_nodes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(num_nodes >= 0, LGK_ERR_BAD_ARG, "lgk_gcn_norm: negative size");
  if (num_nodes == 0) return LGK_OK;
  LGK_REQUIRE(rowptr && dinv, LGK_ERR_BAD_ARG, "lgk_gcn_norm: null pointer");
  gcn_norm_kernel<<<unsigned(ceil_div(num_nodes, 256)), 256, 0, stream>>>(rowptr, col, ew, fill, add_self_loops, dinv,
                                                                          selfw_out, num_nodes);
  LGK_CHECK_LAUNCH("gcn_norm_kernel");
  return LGK_OK;
}
