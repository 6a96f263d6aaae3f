// SortAggregation read-out (torch_geometric.nn.aggr.SortAggregation(k); reference models/drgnet.py:37,59;
// SURVEY.md 8f row 2): per graph, order the nodes by their LAST feature channel, descending and stable, keep the first
// k rows, zero-pad graphs with fewer than k nodes, flatten to [B, k*F].
//
// PyG densifies the batch to [B, n_max, F], sorts and gathers it.  Here one CTA owns one graph slab (a contiguous
// [n_g, F] block of x): the keys go to shared memory once, every node finds its rank by counting the nodes that
// precede it (n_g^2 / 256 comparisons per thread: lesion graphs have tens to hundreds of nodes; slabs beyond the
// shared-memory budget re-read the keys through L1/L2), the nodes ranked below k publish their index, and the CTA
// copies the selected rows.  No atomics, no global scratch: bit-reproducible.  The backward is the transposed copy
// through the saved selection (every node is selected at most once).
#include "common.cuh"

namespace lgk {
namespace {

constexpr int kThreads = 256;
constexpr int kKeyCap = 8192;  // keys held in shared memory (32 KB)

// torch.sort(descending=True, stable=True) order: NaN first, then larger values, ties by position
__device__ __forceinline__ bool precedes(float a, int ia, float b, int ib) {
  const bool na = a != a, nb = b != b;
  if (na || nb) return (na && !nb) || (na && nb && ia < ib);
  return a > b || (a == b && ia < ib);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) sort_pool_fwd_kernel(const T* __restrict__ x, const int* __restrict__ ptr,
                                                                 T* __restrict__ out, int* __restrict__ sel, int F, int k) {
  __shared__ float s_key[kKeyCap];
  extern __shared__ int s_sel[];  // [k]
  const int g = blockIdx.x;
  const int lo = ptr[g], n = ptr[g + 1] - lo;
  const T* key = x + int64_t(lo) * F + (F - 1);
  const bool cached = n <= kKeyCap;
  for (int t = threadIdx.x; t < k; t += kThreads) s_sel[t] = -1;
  if (cached)
    for (int t = threadIdx.x; t < n; t += kThreads) s_key[t] = Vec4<T>::load1(key, int64_t(t) * F);
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += kThreads) {
    const float a = cached ? s_key[i] : Vec4<T>::load1(key, int64_t(i) * F);
    int rank = 0;
    if (cached) {
      for (int j = 0; j < n; ++j) rank += precedes(s_key[j], j, a, i) ? 1 : 0;
    } else {
      for (int j = 0; j < n && rank < k; ++j) rank += precedes(Vec4<T>::load1(key, int64_t(j) * F), j, a, i) ? 1 : 0;
    }
    if (rank < k) s_sel[rank] = lo + i;  // ranks are a permutation: one writer per slot
  }
  __syncthreads();
  for (int t = threadIdx.x; t < k; t += kThreads) sel[int64_t(g) * k + t] = s_sel[t];
  T* o = out + int64_t(g) * k * F;
  for (int r = 0; r < k; ++r) {
    const int node = s_sel[r];
    if (node >= 0) {
      const T* src = x + int64_t(node) * F;
      for (int f = threadIdx.x; f < F; f += kThreads) o[int64_t(r) * F + f] = src[f];
    } else {
      for (int f = threadIdx.x; f < F; f += kThreads) Vec4<T>::store1(o, int64_t(r) * F + f, 0.f);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) sort_pool_bwd_kernel(const T* __restrict__ dy, const int* __restrict__ sel,
                                                                 T* __restrict__ dx, int64_t slots, int F) {
  const int64_t slot = int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5);
  if (slot >= slots) return;
  const int node = sel[slot];
  if (node < 0) return;
  const int lane = threadIdx.x & 31;
  for (int f = lane; f < F; f += 32) dx[int64_t(node) * F + f] = dy[slot * F + f];
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int lgk_sort_pool_fwd(const void* x, const int32_t* ptr, void* out, int32_t* sel, int64_t num_graphs, int64_t feat,
                                 int64_t k, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(num_graphs >= 0 && feat > 0 && k > 0, LGK_ERR_BAD_ARG, "lgk_sort_pool_fwd: bad sizes");
  LGK_REQUIRE(k <= 2048 && feat < (int64_t(1) << 30), LGK_ERR_UNSUPPORTED, "lgk_sort_pool_fwd: k or feat out of range");
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_BAD_ARG, "lgk_sort_pool_fwd: bad dtype %d", dtype);
  if (num_graphs == 0) return LGK_OK;
  LGK_REQUIRE(x && ptr && out && sel, LGK_ERR_BAD_ARG, "lgk_sort_pool_fwd: null pointer");
  const size_t smem = size_t(k) * sizeof(int);
  if (dtype == LGK_F32)
    sort_pool_fwd_kernel<float><<<unsigned(num_graphs), kThreads, smem, stream>>>(
        static_cast<const float*>(x), ptr, static_cast<float*>(out), sel, int(feat), int(k));
  else
    sort_pool_fwd_kernel<__nv_bfloat16><<<unsigned(num_graphs), kThreads, smem, stream>>>(
        static_cast<const __nv_bfloat16*>(x), ptr, static_cast<__nv_bfloat16*>(out), sel, int(feat), int(k));
  LGK_CHECK_LAUNCH("sort_pool_fwd_kernel");
  return LGK_OK;
}

extern "C" int lgk_sort_pool_bwd(const void* dy, const int32_t* sel, void* dx, int64_t num_nodes, int64_t num_graphs,
                                 int64_t feat, int64_t k, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(num_nodes >= 0 && num_graphs >= 0 && feat > 0 && k > 0, LGK_ERR_BAD_ARG, "lgk_sort_pool_bwd: bad sizes");
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_BAD_ARG, "lgk_sort_pool_bwd: bad dtype %d", dtype);
  if (num_nodes == 0) return LGK_OK;
  LGK_REQUIRE(dx, LGK_ERR_BAD_ARG, "lgk_sort_pool_bwd: null pointer");
  const size_t esz = dtype == LGK_F32 ? 4 : 2;
  cudaError_t e = cudaMemsetAsync(dx, 0, size_t(num_nodes) * size_t(feat) * esz, stream);  // unselected nodes: zero gradient
  LGK_REQUIRE(e == cudaSuccess, LGK_ERR_LAUNCH, "lgk_sort_pool_bwd: memset failed: %s", cudaGetErrorString(e));
  const int64_t slots = num_graphs * k;
  if (slots == 0) return LGK_OK;
  LGK_REQUIRE(dy && sel, LGK_ERR_BAD_ARG, "lgk_sort_pool_bwd: null pointer");
  const unsigned grid = unsigned(ceil_div(slots, kThreads / 32));
  if (dtype == LGK_F32)
    sort_pool_bwd_kernel<float><<<grid, kThreads, 0, stream>>>(static_cast<const float*>(dy), sel, static_cast<float*>(dx), slots, int(feat));
  else
    sort_pool_bwd_kernel<__nv_bfloat16><<<grid, kThreads, 0, stream>>>(static_cast<const __nv_bfloat16*>(dy), sel,
                                                                     static_cast<__nv_bfloat16*>(dx), slots, int(feat));
  LGK_CHECK_LAUNCH("sort_pool_bwd_kernel");
  return LGK_OK;
}
